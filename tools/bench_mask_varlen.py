"""Device time of the attn_mask and variable-length operators (development tool; whole operator, quantisation included).
    [SAGE_B200_LIB=...] python tools/bench_mask_varlen.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S  # noqa: E402


def t(fn, reps=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


name = os.path.basename(os.environ.get("SAGE_B200_LIB", "prod"))
dev = "cuda:0"
torch.manual_seed(0)
for d in (128, 64):
    b, h, n = 2, 16, 4096
    q, k, v = (torch.randn(b, h, n, d, device=dev, dtype=torch.float16) for _ in range(3))
    mask = torch.rand(b, 1, n, n, device=dev) < 0.7
    ms = t(lambda: S.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=mask))
    print(f"{name:20s} attn_mask (bool) B{b} H{h} N{n} D{d}: {ms:.3f} ms = {4.0 * b * h * n * n * d / ms * 1e-9:.0f} TOPS")
    lens = [4096, 1000, 3000, 2048, 517, 4096, 2500, 3500]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
    tot = sum(lens)
    qv, kv, vv = (torch.randn(tot, h, d, device=dev, dtype=torch.float16) for _ in range(3))
    ms = t(lambda: S.sageattn_varlen(qv, kv, vv, cu, cu, max(lens), max(lens)))
    fl = sum(4.0 * h * L * L * d for L in lens)
    print(f"{name:20s} sageattn_varlen {len(lens)} sequences, {tot} tokens, H{h} D{d}: {ms:.3f} ms = {fl / ms * 1e-9:.0f} TOPS")
