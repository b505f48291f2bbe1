"""attn_mask (the mask half): sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=...), reference
/root/reference/sageattention/core.py:160-361 and triton/attn_qk_int8_per_thread.py:22-166.

(a) the masked sm_100a kernels against the reference's own Triton kernel run on the same GPU with the SAME int8
    codes, scales and mask (the Triton kernel accumulates P.V in fp16: cos >= 0.9999, rel-L1 <= 1e-2);
(b) the operator end to end against fp32 attention with the reference's mask semantics (bool: True = attend; a
    mask of q's dtype is added to the BASE-2 logits as it is): cos >= 0.999, rel-L1 <= 1.5e-2, |dLSE| <= 5e-2;
(c) edge cases: rows with every key masked (zeros, no NaN), broadcast shapes, both layouts, kv lengths that are not
    a multiple of 128, masks used in place vs padded copies, head_dim padding, error conventions.
"""
import math

import pytest
import torch

import sageattention_b200 as sb
from sageattention_b200 import _lib, core
from conftest import accuracy

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LOG2E = 1.4426950408889634


def make_qkv(b, hq, hk, nq, nk, d, dtype, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    q = torch.randn(b, hq, nq, d, generator=g, device=DEV)
    k = torch.randn(b, hk, nk, d, generator=g, device=DEV) + 1.5 * torch.randn(1, hk, 1, d, generator=g, device=DEV)
    v = torch.randn(b, hk, nk, d, generator=g, device=DEV)
    return q.to(dtype), k.to(dtype), v.to(dtype)


def make_mask(kind, shape, dtype, seed, keep=0.6):
    g = torch.Generator(device=DEV).manual_seed(seed)
    keep_m = torch.rand(shape, generator=g, device=DEV) < keep
    keep_m[..., 0] = True                      # no empty rows here (tested separately)
    if kind == "bool":
        return keep_m
    if kind == "neg_inf":
        return torch.zeros(shape, device=DEV, dtype=dtype).masked_fill(~keep_m, float("-inf"))
    bias = torch.randn(shape, generator=g, device=DEV) * 2.0   # finite additive bias + some hard masking
    return bias.masked_fill(~keep_m, -30000.0).to(dtype)


def reference_attention(q, k, v, mask, sm_scale):
    """fp32 attention with the reference's mask semantics (triton/attn_qk_int8_per_thread.py:65-72): the logits are
    in base 2, a bool mask removes keys, a float mask is added to the base-2 logits unscaled."""
    g = q.size(1) // k.size(1)
    s = (q.float() @ k.float().repeat_interleave(g, 1).transpose(2, 3)) * (sm_scale * LOG2E)
    if mask.dtype == torch.bool:
        s = s.masked_fill(~mask, float("-inf"))
    else:
        s = s + mask.float()
    m = s.amax(-1, keepdim=True).clamp_min(-1e30)
    p = torch.exp2(s - m)
    d = p.sum(-1, keepdim=True)
    o = (p / d.clamp_min(1e-38)) @ v.float().repeat_interleave(g, 1)
    lse2 = (torch.log2(d) + m).squeeze(-1)
    return o, lse2 / LOG2E


@pytest.mark.parametrize("kind", ["bool", "float"])
@pytest.mark.parametrize("b,hq,hk,nq,nk,d", [(2, 4, 4, 512, 512, 128), (1, 4, 2, 300, 700, 64), (1, 2, 2, 130, 515, 128)])
def test_masked_kernel_vs_reference_triton_kernel_same_codes(ref_triton, kind, b, hq, hk, nq, nk, d):
    tri = ref_triton("attn_qk_int8_per_thread")
    dtype = torch.float16
    q, k, v = make_qkv(b, hq, hk, nq, nk, d, dtype, seed=nq + d)
    km = k.mean(dim=2)
    q8, qs, k8, ks = sb.per_thread_int8(q, k, km)
    mask = make_mask(kind, (b, hq, nq, nk), dtype, seed=7)
    sm = d ** -0.5
    o = torch.empty_like(q)
    nb, _, _, _, _ = core._dims(q, "HND")
    core._launch(False, q8, k8, v, o, qs, ks, None, None, "HND", False, _lib.GRAN_PER_THREAD, 32, sm, False,
                 mask=core._prepare_mask(mask, q, b, hq, nq, nk))
    ro, _ = tri.forward(q8, k8, v, qs, ks, sm, tensor_layout="HND", attn_mask=mask, output_dtype=dtype)
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all()
    cos, l1, _ = accuracy(o, ro)
    assert cos >= 0.9999 and l1 <= 1e-2, (cos, l1)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("kind,mshape", [("bool", "full"), ("bool", "b1"), ("bool", "11"), ("neg_inf", "full"),
                                         ("float", "1h"), ("bool", "2d")])
@pytest.mark.parametrize("b,hq,hk,nq,nk,d", [(2, 4, 2, 384, 384, 128), (1, 4, 4, 200, 333, 64), (2, 2, 2, 160, 1000, 96)])
def test_masked_operator_vs_fp32_attention(dtype, layout, kind, mshape, b, hq, hk, nq, nk, d):
    q, k, v = make_qkv(b, hq, hk, nq, nk, d, dtype, seed=3 * nq + d)
    shape = {"full": (b, hq, nq, nk), "b1": (b, 1, nq, nk), "11": (1, 1, nq, nk), "1h": (1, hq, nq, nk),
             "2d": (nq, nk)}[mshape]
    mask = make_mask(kind, shape, dtype, seed=11)
    sm = d ** -0.5
    m4 = mask if mask.dim() == 4 else mask[None, None]
    ref, ref_lse = reference_attention(q, k, v, m4.expand(b, hq, nq, nk), sm)
    if layout == "NHD":
        qi, ki, vi = (t.transpose(1, 2).contiguous() for t in (q, k, v))
    else:
        qi, ki, vi = q, k, v
    o, lse = sb.sageattn_qk_int8_pv_fp16_triton(qi, ki, vi, tensor_layout=layout, attn_mask=mask, return_lse=True)
    assert o.shape == qi.shape and o.dtype == dtype
    oh = o.transpose(1, 2) if layout == "NHD" else o
    assert torch.isfinite(oh.float()).all()
    cos, l1, _ = accuracy(oh, ref)
    assert cos >= 0.999 and l1 <= 1.5e-2, (cos, l1)
    assert (lse - ref_lse).abs().max().item() <= 5e-2


def test_rows_with_every_key_masked_give_zeros_and_tiles_of_all_false_cost_nothing_wrong():
    b, h, nq, nk, d = 1, 2, 300, 600, 128
    q, k, v = make_qkv(b, h, h, nq, nk, d, torch.float16, seed=5)
    mask = make_mask("bool", (b, h, nq, nk), torch.float16, seed=6)
    mask[:, :, 17] = False             # a row with nothing to attend to
    mask[:, :, 200:260] = False        # a whole band of such rows
    mask[:, :, :, 128:384] = False     # two whole kv tiles masked for every row
    mask[:, :, 100, :] = False
    mask[:, :, 100, 599] = True        # a single key, in the last (ragged) tile
    o, lse = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=mask, return_lse=True)
    assert torch.isfinite(o.float()).all()
    assert (o[:, :, 17] == 0).all() and (o[:, :, 200:260] == 0).all()
    assert torch.isinf(lse[:, :, 17]).all() and (lse[:, :, 17] < 0).all()
    # one key left: the output row is that value row (to f16 rounding of p = 1)
    assert torch.allclose(o[:, :, 100].float(), v[:, :, 599].float(), atol=2e-3, rtol=2e-3)
    ref, _ = reference_attention(q, k, v, mask, d ** -0.5)
    live = mask.any(-1)
    cos, l1, _ = accuracy(o[live], ref[live])
    assert cos >= 0.999 and l1 <= 1.5e-2, (cos, l1)


def test_mask_in_place_equals_padded_copy_and_no_mask_equals_cuda_entry_point():
    b, h, nq, nk, d = 2, 4, 256, 512, 128      # nk a multiple of 128, contiguous: the mask is read in place
    q, k, v = make_qkv(b, h, h, nq, nk, d, torch.bfloat16, seed=9)
    mask = make_mask("bool", (b, 1, nq, nk), torch.bfloat16, seed=10)
    prepared = core._prepare_mask(mask, q, b, h, nq, nk)
    assert prepared[0].data_ptr() == mask.data_ptr() and prepared[2] == mask.stride(0) and prepared[3] == 0
    o1 = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=mask)
    # the same mask as a strided view (forces the padded copy): same bits out
    big = torch.zeros(b, 1, nq, nk + 8, dtype=torch.bool, device=DEV)
    big[..., 3:3 + nk] = mask
    view = big[..., 3:3 + nk]
    assert core._prepare_mask(view, q, b, h, nq, nk)[0].data_ptr() != view.data_ptr()
    o2 = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=view)
    assert torch.equal(o1, o2)
    # all-True mask == no mask, up to the different softmax path (float-domain max instead of the integer one)
    o3 = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=torch.ones(1, 1, nq, nk, dtype=torch.bool, device=DEV))
    o4 = sb.sageattn_qk_int8_pv_fp16_triton(q, k, v)
    o5 = sb.sageattn_qk_int8_pv_fp16_cuda(q, k, v)
    assert torch.equal(o4, o5)
    cos, l1, _ = accuracy(o3, o4)
    assert cos >= 0.99999 and l1 <= 2e-3, (cos, l1)


def test_mask_error_conventions():
    q, k, v = make_qkv(1, 2, 2, 128, 128, 64, torch.float16, seed=1)
    mask = torch.ones(1, 1, 128, 128, dtype=torch.bool, device=DEV)
    with pytest.raises(AssertionError):      # core.py:303 "Mask should be None for causal attention."
        sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, is_causal=True, attn_mask=mask)
    with pytest.raises(AssertionError):      # core.py:250: bool or q's dtype
        sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=mask.float())
    with pytest.raises(AssertionError):      # core.py:314-315: not broadcastable
        sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, attn_mask=torch.ones(1, 1, 64, 128, dtype=torch.bool, device=DEV))
    with pytest.raises(ValueError):
        sb.sageattn_qk_int8_pv_fp16_triton(q, k, v, quantization_backend="numpy")
    # the C ABI refuses a mask together with causal, and misaligned rows
    o = torch.empty_like(q)
    q8, qs, k8, ks = sb.per_thread_int8(q, k, None)
    with pytest.raises(ValueError):
        core._launch(False, q8, k8, v, o, qs, ks, None, None, "HND", True, 3, 32, 0.125, False,
                     mask=core._prepare_mask(mask, q, 1, 2, 128, 128))
    with pytest.raises(ValueError):
        core._launch(False, q8, k8, v, o, qs, ks, None, None, "HND", False, 3, 32, 0.125, False,
                     mask=(mask, 1, 0, 0, 100))
