import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from wav2vec2forbrain_b200 import ops
import gemm_bench as gb
M = 3776
for sk in (False, True):
    ops.STREAM_K = sk
    print("--- stream-K", sk, flush=True)
    for bn in (128, 256):
        gb.bench("out proj (N=1024,K=1024)", M, 1024, 1024, True, bn=bn)
        gb.bench("dgrad 1024 (B mn)", M, 1024, 1024, True, b_mn=True, bn=bn)
        gb.bench("qk proj (N=2048,K=1024)", M, 2048, 1024, True, bn=bn)
        gb.bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, True, bn=bn)
        gb.bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, True, bn=bn)
        gb.bench("wgrad 1024x1024 (sk2)", 1024, 1024, M, True, a_mn=True, b_mn=True, split_k=2, bn=bn)
        gb.bench("wgrad 2048x1024", 2048, 1024, M, True, a_mn=True, b_mn=True, bn=bn)
        gb.bench("wgrad ffn1 4096x1024", 4096, 1024, M, True, a_mn=True, b_mn=True, bn=bn)
        gb.bench("base ffn2 (M=7552,N=768,K=3072)", 7552, 768, 3072, True, bn=bn)
        gb.bench("base out (M=7552,N=768,K=768)", 7552, 768, 768, True, bn=bn)
