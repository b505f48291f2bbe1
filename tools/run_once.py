"""Run the attention operator a few times on one shape (target for ncu captures).
    [SAGE_B200_LIB=...] python tools/run_once.py [f8|f16] [head_dim] [seq] [batch] [heads] [causal]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S

kind = sys.argv[1] if len(sys.argv) > 1 else "f8"
D = int(sys.argv[2]) if len(sys.argv) > 2 else 128
N = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
B = int(sys.argv[4]) if len(sys.argv) > 4 else 1
H = int(sys.argv[5]) if len(sys.argv) > 5 else 8
causal = bool(int(sys.argv[6])) if len(sys.argv) > 6 else False
torch.manual_seed(0)
q, k, v = (torch.randn(B, H, N, D, device="cuda", dtype=torch.float16) for _ in range(3))
fn = S.sageattn_qk_int8_pv_fp8_cuda if kind == "f8" else S.sageattn_qk_int8_pv_fp16_cuda
for _ in range(3):
    o = fn(q, k, v, tensor_layout="HND", is_causal=causal)
torch.cuda.synchronize()
print("ok", kind, D, N, B, H, causal, float(o.float().abs().mean()))
