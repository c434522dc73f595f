"""Host-side issue cost of one training step WITHOUT a GPU: the product's Python path (models / engine / ops / ctypes
argument packing / torch allocator calls) runs unchanged on CPU tensors, the C ABI is replaced by entry points that return 0
at once.  Numbers are garbage (nothing is computed); what is measured is the Python + ctypes time per step, which is what
bounds the step when many ranks share one host.  Usage:
    python scripts/host_null.py [base|conformer] [B] [--profile]
Not a test and not a bench: a development tool for this container, where no GPU exists."""
import cProfile
import ctypes
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from wav2vec2forbrain_b200 import _lib


class _Fn:
    """stands for a C entry point: accepts anything, returns 0; counts calls"""
    __slots__ = ("name", "calls", "restype", "argtypes")

    def __init__(self, name):
        self.name, self.calls, self.restype, self.argtypes = name, 0, ctypes.c_int, None

    def __call__(self, *a):
        self.calls += 1
        if self.name == "b2t_last_error":
            return b""
        if self.name.endswith("_bytes") or self.name.endswith("_floats"):
            return 1 << 20
        if self.name == "b2t_gru_max_active_clusters":
            return 7
        return 0


class NullLib:
    def __init__(self):
        self._fns = {}

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        f = self._fns.get(name)
        if f is None:
            f = self._fns[name] = _Fn(name)
        return f

    def counts(self):
        return {k: v.calls for k, v in sorted(self._fns.items()) if v.calls}


def install():
    null = NullLib()
    _lib._lib = null
    _lib._device_ok = True
    _lib.raw_stream = lambda: 0
    _lib.stream = lambda: ctypes.c_void_p(0)
    from wav2vec2forbrain_b200 import models, ops
    ops.stream = _lib.stream
    models._EMULATED = True
    # data movement that would be an asynchronous kernel on the GPU must not be paid as real CPU work here
    _empty = torch.empty
    torch.zeros = lambda *shape, **kw: _empty(*shape, **{k: v for k, v in kw.items() if k != "out"})
    torch.zeros_like = torch.empty_like
    torch.Tensor.zero_ = lambda self: self
    _stack = torch.stack
    torch.stack = lambda ts, dim=0: _empty((len(ts),) + tuple(ts[0].shape), dtype=ts[0].dtype) if dim == 0 else _stack(ts, dim)
    # (run with python -O: ops.py asserts is_cuda in one place; CPU tensors stand for device tensors here)
    return null


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "conformer"
    B = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else (32 if wl == "conformer" else 64)
    null = install()
    import bench
    from wav2vec2forbrain_b200 import models
    torch.nn.Module.cuda = lambda self, *a, **k: self
    t0 = time.perf_counter()
    model, _ = bench.build_model(wl, "bf16x3", optimizer="none")
    print(f"model built in {time.perf_counter() - t0:.1f} s")
    hb = bench.make_host_batch(B, pin=False)
    batch = bench.to_device_batch(models, hb, torch.device("cpu"))

    def step():
        model.zero_grad(set_to_none=True)
        out = model.forward(batch)
        out.loss.backward()

    for _ in range(3):
        step()
    ts = []
    for _ in range(7):
        t0 = time.perf_counter()
        step()
        ts.append(1e3 * (time.perf_counter() - t0))
    ts.sort()
    print(f"issue time per step (forward + backward, no optimizer): min {ts[0]:.1f} median {ts[len(ts) // 2]:.1f} max {ts[-1]:.1f} ms")
    before = dict(null.counts())
    step()
    after = null.counts()
    per = {k: after[k] - before.get(k, 0) for k in after if after[k] - before.get(k, 0)}
    print("entry-point calls per step:", sum(per.values()), per)
    if "--profile" in sys.argv:
        pr = cProfile.Profile()
        pr.enable()
        step()
        pr.disable()
        st = pstats.Stats(pr)
        st.sort_stats("tottime").print_stats(45)
        st.sort_stats("cumulative").print_stats(45)


if __name__ == "__main__":
    main()
