"""development: the statistics kernel alone, for an ncu capture (python tools/dbg_stats_ncu.py)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sageattention_b200 import quant
for shape in ((4, 32, 4096, 128), (1, 40, 75600, 128)):
    k = torch.randn(shape, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        km = quant.seq_mean(k, "HND")
    torch.cuda.synchronize()
