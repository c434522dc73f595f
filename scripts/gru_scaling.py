import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from wav2vec2forbrain_b200 import ops, _lib
_lib.ensure_device()
print("max active 16-CTA clusters:", _lib.lib().b2t_gru_max_active_clusters(512))
import kernel_bench as kb
for B in (16, 32, 48, 64, 96, 128):
    kb.gru(B)
