"""b2t_ctc_fwd_bwd_static: the allocation-free variant of the CTC entry point (scratch in a static device array) against
b2t_ctc_fwd_bwd with a caller-owned workspace, through the C ABI, on two calls in a row.  Collected last on
purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-6      # the same kernel runs behind both entry points: only the scratch address differs


def _call(lib, static, logits, targets, input_lens, target_lens):
    from wav2vec2forbrain_b200._lib import c_int, c_int64, check, ptr, stream
    B, T, V = logits.shape
    S = targets.shape[1]
    nll = torch.empty(B, dtype=torch.float32, device="cuda")
    loss = torch.zeros((), dtype=torch.float32, device="cuda")
    grad = torch.empty_like(logits)
    common = (ptr(logits), c_int64(logits.stride(1)), c_int64(logits.stride(0)), ptr(targets), c_int(S), ptr(input_lens),
              ptr(target_lens))
    tail = (ptr(nll), ptr(loss), ptr(grad), c_int(B), c_int(T), c_int(V), c_int(1), stream())
    if static:
        rc = lib.b2t_ctc_fwd_bwd_static(*common, *tail)
        check(rc, "b2t_ctc_fwd_bwd_static")
    else:
        lib.b2t_ctc_workspace_floats.restype = ctypes.c_int64
        ws = torch.empty(lib.b2t_ctc_workspace_floats(c_int(B), c_int(T), c_int(S)), dtype=torch.float32, device="cuda")
        check(lib.b2t_ctc_fwd_bwd(*common, ptr(ws), *tail), "b2t_ctc_fwd_bwd")
    torch.cuda.synchronize()
    return loss, nll, grad


def _case(B, T, S, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    logits = torch.randn(B, T, 32, generator=g).cuda()
    target_lens = torch.randint(0, S + 1, (B,), generator=g, dtype=torch.int64)
    target_lens[0] = S
    targets = torch.randint(1, 32, (B, S), generator=g, dtype=torch.int64)
    input_lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g, dtype=torch.int32)
    input_lens[0] = T
    return logits, targets.cuda(), input_lens.cuda(), target_lens.cuda()


def _close(a, b):
    a, b = a.double(), b.double()
    return (a - b).norm().item() <= TOL * b.norm().item() + 1e-12


def test_static_scratch_ctc_matches_the_workspace_entry_point_twice():
    from wav2vec2forbrain_b200 import _lib
    _lib.ensure_device()
    lib = _lib.lib()
    lib.b2t_ctc_static_max_floats.restype = ctypes.c_int64
    assert lib.b2t_ctc_static_max_floats() == 2 * 1024 * 1024
    # two different problems back to back: the second one finds the first one's lattice in the shared scratch
    for i, (B, T, S) in enumerate([(64, 118, 40), (37, 90, 57)]):
        case = _case(B, T, S, seed=100 + i)
        ref = _call(lib, False, *case)
        if i == 1:
            free_before = torch.cuda.mem_get_info()[0]
        got = _call(lib, True, *case)
        if i == 1:      # the entry point itself allocates nothing on the device (the outputs above are the caller's)
            assert torch.cuda.mem_get_info()[0] >= free_before - (B * T * 32 + B + 1) * 4 - (2 << 20)
        for r, o, what in zip(ref, got, ("loss", "nll", "grad")):
            assert torch.isfinite(o).all(), what
            assert _close(o, r), (what, i)
    # and the same problem again gives the same answer (nothing in the scratch has to be cleared)
    again = _call(lib, True, *case)
    assert torch.equal(again[2], got[2]) and torch.equal(again[1], got[1])


def test_static_scratch_ctc_rejects_sizes_beyond_its_maximum():
    from wav2vec2forbrain_b200 import _lib
    _lib.ensure_device()
    lib = _lib.lib()
    case = _case(300, 118, 40, seed=7)      # 300 * 118 * 128 floats > 2 Mi
    with pytest.raises(RuntimeError, match="static scratch holds"):
        _call(lib, True, *case)
    ok = _call(lib, False, *case)            # the workspace entry point takes the same problem
    assert torch.isfinite(ok[0])
