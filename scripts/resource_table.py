"""Per-kernel registers / stack (spill) / static shared memory of the built library, from cuobjdump --dump-resource-usage.
Usage: python scripts/resource_table.py > profiles/resources_rNN.txt   (no GPU needed)"""
import os
import re
import subprocess
import sys

so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "wav2vec2forbrain_b200", "csrc", "libb2t_sm100.so")
txt = subprocess.run(["cuobjdump", "--dump-resource-usage", so], text=True, capture_output=True).stdout
rows, ident = [], ""
lines = txt.splitlines()
for i, l in enumerate(lines):
    if l.startswith("identifier"):
        ident = l.split("=")[1].strip()
    m = re.match(r"\s*Function (\S+):$", l)
    if m:
        d = dict(kv.split(":") for kv in lines[i + 1].split() if ":" in kv)
        rows.append((ident, m.group(1), d))
dem = subprocess.run(["c++filt"], input="\n".join(r[1] for r in rows), text=True, capture_output=True).stdout.splitlines()


def short(s):
    s = re.sub(r"\(anonymous namespace\)::", "", s).replace("void ", "")
    depth = 0
    for j in range(len(s) - 1, -1, -1):          # strip the trailing parameter list
        if s[j] == ")":
            depth += 1
        elif s[j] == "(":
            depth -= 1
            if depth == 0:
                return s[:j]
    return s


print(f"{'file':<20}{'REG':>4}{'STACK':>6}{'SHARED':>8}{'LOCAL':>6}  kernel")
n_stack = 0
for (ident, _, d), dn in sorted(zip(rows, dem), key=lambda x: (x[0][0], x[1])):
    st = int(d.get("STACK", 0))
    n_stack += st > 0
    print(f"{ident:<20}{d.get('REG', '?'):>4}{st:>6}{d.get('SHARED', '?'):>8}{d.get('LOCAL', '0'):>6}  {short(dn)}")
print(f"\n{len(rows)} kernels, {n_stack} with a stack frame (spills or local arrays)", file=sys.stdout)
