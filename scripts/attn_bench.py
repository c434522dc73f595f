import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/scripts')
import kernel_bench as kb
kb.attn(64, 12); kb.attn(32, 16)
