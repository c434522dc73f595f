"""Helpers shared by the oracle/golden tests and the GPU parity tests."""
import os

import torch

from oracle import b2t_oracle as O

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_case(name):
    rec = torch.load(os.path.join(GOLDEN_DIR, name + ".pt"), weights_only=False)
    be = O.BrainEncoderCfg(**rec["be"])
    w2v = O.W2VCfg(**rec["w2v"])
    sd = O.make_state_dict(be, w2v, seed=rec["seed_w"])
    batch = O.make_batch(rec["B"], rec["T"], rec["S"], seed=rec["seed_x"], ragged=rec["ragged"])
    return rec, be, w2v, sd, batch


def rel_l2(a, b):
    a = a.double().flatten()
    b = b.double().flatten()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def golden_grad(rec, key):
    """Return (reference tensor, slicer) for parameter `key` from a full-tensor golden record."""
    g = rec["grads"]
    if key in g:
        return g[key], (lambda t: t)
    return g[key + "::11"], (lambda t: t.flatten()[::11])
