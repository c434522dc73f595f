"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one training step
(the launches between two consecutive CTC kernels) grouped by kernel, with count, total time and share.
usage: python profiles/summarize_launches.py launches.csv [step_index] [marker-regex]"""
import collections
import csv
import re
import sys


def short(name):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    m = re.match(r"(?:void )?([\w:]+)(<[^(]*>)?\(", name)
    if not m:
        return name[:80]
    base, targs = m.group(1), m.group(2) or ""
    if base.startswith("at::") or "at::native" in name:
        inner = re.search(r"at::native::(\w+)", name)
        return "torch:" + (inner.group(1) if inner else base)[:60]
    if len(targs) > 40:
        targs = targs[:40] + "...>"
    return base + targs


def main():
    path = sys.argv[1]
    step = int(sys.argv[2]) if len(sys.argv) > 2 else -2
    marker = sys.argv[3] if len(sys.argv) > 3 else "ctc_"
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    marks = [i for i, r in enumerate(rows) if re.search(marker, r["Kernel Name"])]
    if len(marks) < 2:
        sel = rows
        print(f"# no step marker found; summarising all {len(rows)} launches")
    else:
        lo, hi = marks[step], marks[step + 1] if step + 1 != 0 else len(rows)
        sel = rows[lo:hi]
        print(f"# step between marker launches {lo} and {hi} ({len(sel)} launches; {len(marks)} markers in capture)")
    agg = collections.OrderedDict()
    for r in sel:
        k = short(r["Kernel Name"])
        c, t = agg.get(k, (0, 0.0))
        agg[k] = (c + 1, t + float(r["Metric Value"]) / 1e3)
    total = sum(t for _, t in agg.values())
    print(f"# total device time of the step (serialised, cold-cache): {total / 1e3:.3f} ms")
    print(f"{'kernel':70s} {'n':>5s} {'us':>10s} {'share':>7s}")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:70s} {c:5d} {t:10.1f} {100 * t / total:6.1f}%")


if __name__ == "__main__":
    main()
