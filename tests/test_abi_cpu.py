"""The C-ABI boundary without a GPU: libb2t_sm100.so loads, exports every function include/b2t_sm100.h
declares, and the ctypes mirror of b2t_gemm_args has the C compiler's layout.  No compute calls."""
import ctypes
import os
import re
import subprocess
import sys
import tempfile

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(REPO, "include", "b2t_sm100.h")
LIB = os.path.join(REPO, "wav2vec2forbrain_b200", "csrc", "libb2t_sm100.so")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2t_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_path_families():
    names = declared_functions()
    for fam in ("b2t_gemm", "b2t_smooth_fwd", "b2t_layernorm_fwd", "b2t_layernorm_bwd", "b2t_attention_fwd", "b2t_attention_bwd",
                "b2t_gru_fwd", "b2t_gru_bwd", "b2t_layernorm_rope_fwd", "b2t_glu_dwconv_fwd", "b2t_bn_silu_fwd", "b2t_ctc_fwd_bwd", "b2t_greedy_decode", "b2t_abi_version", "b2t_last_error"):
        assert fam in names, fam
    assert len(names) >= 30


@pytest.mark.skipif(not os.path.exists(LIB), reason="libb2t_sm100.so not built (python -c 'import __graft_entry__ as g; g.build()')")
def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(LIB)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in include/b2t_sm100.h but not exported: {missing}"
    lib.b2t_abi_version.restype = ctypes.c_int
    assert lib.b2t_abi_version() == 1


def test_gemm_args_layout_matches_the_c_compiler():
    """sizeof / offsetof of b2t_gemm_args from gcc against the ctypes Structure the host code fills."""
    sys.path.insert(0, REPO)
    from wav2vec2forbrain_b200._lib import GemmArgs
    fields = [f[0] for f in GemmArgs._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {',
            '  printf("sizeof %zu\\n", sizeof(b2t_gemm_args));']
    prog += [f'  printf("{f} %zu\\n", offsetof(b2t_gemm_args, {f}));' for f in fields]
    prog += ['  return 0;', '}']
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "layout.c")
        exe = os.path.join(d, "layout")
        with open(c, "w") as fh:
            fh.write("\n".join(prog))
        subprocess.run(["gcc", "-std=c99", "-o", exe, c], check=True)
        out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()
    got = dict(zip(out[0::2], map(int, out[1::2])))
    assert got["sizeof"] == ctypes.sizeof(GemmArgs)
    for f in fields:
        assert got[f] == getattr(GemmArgs, f).offset, f


def test_every_exported_entry_point_is_declared():
    """the other direction: no extern "C" entry point in csrc/ is missing from the header"""
    import glob
    defined = set()
    for path in glob.glob(os.path.join(REPO, "wav2vec2forbrain_b200", "csrc", "*.cu")):
        defined |= set(re.findall(r'extern "C"[^(]*?\b(b2t_[a-z0-9_]+)\s*\(', open(path).read()))
    undeclared = sorted(defined - set(declared_functions()))
    assert not undeclared, undeclared
