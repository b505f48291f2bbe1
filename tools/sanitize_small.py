"""Small instances of every path added in round 2, for `compute-sanitizer --tool memcheck` (development tool):
attn_mask kernels (bool / additive, ragged kv, padded copy), packed variable-length quantiser + attention (partial last
tiles stored row by row, a key-less sequence, an empty query sequence), the Q/K prologue that also writes the f16 V copy,
sage_seq_mean_static, the single-call operator (both P.V flavours, folded head_dim, NHD), a capped grid."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import sageattention_b200 as sb  # noqa: E402
from sageattention_b200 import _lib, core  # noqa: E402
from sageattention_b200.quant import _dims, _dtype_code, _stream  # noqa: E402

dev = "cuda:0"
torch.manual_seed(0)


def qkv(b, hq, hk, nq, nk, d, dt):
    return (torch.randn(b, hq, nq, d, device=dev, dtype=dt), torch.randn(b, hk, nk, d, device=dev, dtype=dt),
            torch.randn(b, hk, nk, d, device=dev, dtype=dt))


# single-call operator
for fn in (sb.sageattn, sb.sageattn_qk_int8_pv_fp16_cuda):
    for dt, d, layout, causal in ((torch.bfloat16, 128, "HND", False), (torch.float16, 64, "NHD", True), (torch.bfloat16, 96, "HND", False)):
        q, k, v = qkv(1, 4, 2, 300, 333, d, dt)
        if causal:
            q, k, v = qkv(1, 4, 2, 300, 300, d, dt)
        if layout == "NHD":
            q, k, v = (t.transpose(1, 2).contiguous() for t in (q, k, v))
        o, lse = fn(q, k, v, tensor_layout=layout, is_causal=causal, return_lse=True)
        assert torch.isfinite(o.float()).all()
# capped grid
core.GRID_LIMIT = 3
q, k, v = qkv(1, 4, 4, 700, 700, 128, torch.float16)
sb.sageattn(q, k, v)
core.GRID_LIMIT = None
# masks
for d in (64, 128):
    q, k, v = qkv(2, 2, 2, 200, 333, d, torch.float16)
    m = torch.rand(2, 1, 200, 333, device=dev) < 0.5
    m[..., 0] = True
    m[:, :, 5] = False
    o = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=m)
    assert torch.isfinite(o.float()).all()
    o = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=torch.zeros(1, 2, 200, 333, device=dev, dtype=torch.float16))
    assert torch.isfinite(o.float()).all()
# varlen
for d, dt in ((128, torch.float16), (64, torch.bfloat16)):
    lens_q, lens_k = [130, 0, 257, 1, 64], [200, 64, 257, 0, 100]
    cq = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32, device=dev)
    ck = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32, device=dev)
    q = torch.randn(sum(lens_q), 4, d, device=dev, dtype=dt)
    k = torch.randn(sum(lens_k), 2, d, device=dev, dtype=dt)
    v = torch.randn(sum(lens_k), 2, d, device=dev, dtype=dt)
    o = sb.sageattn_varlen(q, k, v, cq, ck, max(lens_q), max(lens_k))
    assert torch.isfinite(o.float()).all()
    o = sb.sageattn_varlen(q, q[:, :2].contiguous(), q[:, :2].contiguous(), cq, cq, max(lens_q), max(lens_q), is_causal=True)
    assert torch.isfinite(o.float()).all()
# static-scratch mean
x = torch.randn(2, 3, 1000, 64, device=dev, dtype=torch.bfloat16)
lib = _lib.load()
b, h, n, d, x4 = _dims(x, "HND")
out = torch.empty(b, h, d, device=dev, dtype=x.dtype)
_lib.check(lib.sage_seq_mean_static(x4, out.data_ptr(), b, h, n, d, _dtype_code(x), _stream(x)))
torch.cuda.synchronize()
assert torch.equal(out, sb.seq_mean(x))
print("sanitize_small: ok")
