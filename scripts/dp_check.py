"""2-rank check of the data-parallel path on GPUs: with identical batches and dropout seeds on both ranks the averaged
gradients (early sink inside the encoder node + hook buckets) must equal the local gradients of the same step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from wav2vec2forbrain_b200 import dist as bdist, models

rank, local_rank, world = bdist.env_rank_world()
bdist.init_process_group("nccl")
torch.cuda.set_device(local_rank)
wl = sys.argv[1] if len(sys.argv) > 1 else "conformer"
model, opt = bench.build_model(wl, "bf16x3")
dp = bdist.DataParallel(model)
rt = model.runtime
rt.rank = 0
hb = bench.make_host_batch(8, seed=77)
batch = bench.to_device_batch(models, hb, torch.device("cuda", local_rank))

def grads(step0):
    rt.step = step0
    if hasattr(rt, "layerdrop_gen"):
        rt.layerdrop_gen.manual_seed(123)
    opt.zero_grad(set_to_none=True)
    out = model.forward(batch)
    out.loss.backward()
    dp.finish()
    torch.cuda.synchronize()
    return {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}, out.loss.item()

g_dp, l_dp = grads(0)
dp.buckets.enabled = False      # same schedule, collective switched off: the local gradients of the same step
g_loc, l_loc = grads(0)
worst, wk = 0.0, None
gmax = max(g.double().norm().item() for g in g_loc.values())
for k, g in g_loc.items():
    # absolute floor: gradients that are mathematically zero (k_proj.bias) carry only atomics-order noise
    d = (g_dp[k].double() - g.double()).norm().item() / (g.double().norm().item() + 1e-5 * gmax)
    if d > worst:
        worst, wk = d, k
print(f"rank {rank}: loss dp {l_dp:.6f} local {l_loc:.6f}; {len(g_loc)} gradients, worst rel diff {worst:.2e} at {wk} (|g|max {gmax:.3e})", flush=True)
assert set(g_dp) == set(g_loc)
assert worst < 1e-3, (worst, wk)
dist.barrier()
dist.destroy_process_group()
