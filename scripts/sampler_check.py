"""Does the in-process NVML clock sampler (or host run-ahead) slow the unsynchronised step loop?  Same process, same
model: 20-step loops with the sampler on / off, with and without the weight-gradient stream."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from wav2vec2forbrain_b200 import models, engine
wl = sys.argv[1] if len(sys.argv) > 1 else "conformer"
model, opt = bench.build_model(wl, "bf16x3")
hb = bench.make_host_batch(bench.PER_GPU_BATCH[wl], pin=True)
batch = bench.to_device_batch(models, hb, torch.device("cuda"))
def step():
    opt.zero_grad(set_to_none=True)
    out = model.forward(batch)
    out.loss.backward()
    opt.step()
    return out
for _ in range(5): step()
torch.cuda.synchronize()
def loop(n=20, sync_each=False):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); t0 = time.perf_counter(); e0.record()
    for _ in range(n):
        o = step()
        if sync_each: o.loss.item()
    e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, (t1 - t0) * 1e3 / n
for side in (True, False):
    engine.Engine._side_enabled = side
    for _ in range(3): step()
    for samp in (False, True, False, True):
        s = bench.make_sampler(0) if samp else None
        if s: s.wait_ready(); s.mark()
        ms, host = loop()
        if s: s.stop()
        ms2, _ = loop(sync_each=True)
        print(f"side={side} sampler={samp}: unsynced {ms:.2f} ms/step (host issue {host:.2f}), synced-each-step {ms2:.2f} ms/step, "
              f"reserved {torch.cuda.memory_reserved() / 2**30:.1f} GiB", flush=True)
