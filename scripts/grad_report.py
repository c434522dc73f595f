"""Per-parameter gradient error of the engine against a golden case (diagnostic; GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.engine_util import run_engine_train_step
from tests.golden_util import load_case, rel_l2
from wav2vec2forbrain_b200 import engine
name = sys.argv[1] if len(sys.argv) > 1 else "conformer_large_b2"
rec, _, _, sd, batch = load_case(name)
logits, loss, lens, G, feats, eng, P = run_engine_train_step(engine, rec, sd, batch, "cuda")
print("logits rel", rel_l2(logits.cpu(), rec["logits_train"]), "loss", loss.item(), rec["loss_train"].item())
gmax = max(rec["grad_norm"].values())
out = []
for k, ref in rec["grad_sample"].items():
    v = G[k].cpu()
    got = torch.cat([v.flatten()[:32], v.flatten()[::max(1, v.numel() // 32)][:32]])
    e_s = (got - ref).norm().item() / max(ref.norm().item(), 1e-30)
    e_n = abs(v.double().norm().item() - rec["grad_norm"][k]) / max(rec["grad_norm"][k], 1e-30)
    out.append((max(e_s, e_n), e_s, e_n, ref.norm().item() / gmax, k))
out.sort(reverse=True)
out = [o for o in out if not o[4].endswith('linear_k.bias') and not o[4].endswith('k_proj.bias')]
import statistics
print('median', statistics.median(o[0] for o in out), 'n>5e-4:', sum(o[0] > 5e-4 for o in out), 'of', len(out))
for o in out[:40]:
    print("%.2e sample %.2e norm %.2e  refnorm/gmax %.1e  %s" % o)
