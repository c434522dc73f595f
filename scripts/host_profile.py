"""Host-side cost of one training step (Python + ctypes + allocator), measured with the GPU idle-waiting:
wall time of issuing a step without synchronising, and a cProfile of it."""
import cProfile, pstats, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from wav2vec2forbrain_b200 import models
wl = sys.argv[1] if len(sys.argv) > 1 else "base"
B = int(sys.argv[2]) if len(sys.argv) > 2 else (32 if wl == "conformer" else 64)
model, opt = bench.build_model(wl, "bf16x3")
hb = bench.make_host_batch(B, pin=True)
batch = bench.to_device_batch(models, hb, torch.device("cuda"))
def step():
    opt.zero_grad(set_to_none=True)
    out = model.forward(batch)
    out.loss.backward()
    opt.step()
for _ in range(3): step()
torch.cuda.synchronize()
for i in range(3):
    t0 = time.perf_counter(); step(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"issue {1e3*(t1-t0):.1f} ms, until done {1e3*(t2-t0):.1f} ms")
pr = cProfile.Profile(); pr.enable(); step(); pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
# end-to-end pieces: pinned H2D of one batch, then a full step with the loss read back
dev = torch.device("cuda")
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    bt = bench.to_device_batch(models, hb, dev); torch.cuda.synchronize(); t1 = time.perf_counter()
    opt.zero_grad(set_to_none=True); out = model.forward(bt); torch.cuda.synchronize(); t2 = time.perf_counter()
    out.loss.backward(); torch.cuda.synchronize(); t3 = time.perf_counter()
    opt.step(); torch.cuda.synchronize(); t4 = time.perf_counter()
    l = out.loss.item(); t5 = time.perf_counter()
    print(f"h2d {1e3*(t1-t0):.2f} fwd {1e3*(t2-t1):.2f} bwd {1e3*(t3-t2):.2f} opt {1e3*(t4-t3):.2f} item {1e3*(t5-t4):.2f} ms; pinned={hb['input'].is_pinned()}")
