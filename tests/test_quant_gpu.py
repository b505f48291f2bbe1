"""GPU parity of the pre-pass quantizers (through the C ABI) against
  (a) the reference's own CUDA quantizers compiled from /root/reference (oracle/_ref/_ref_fused.so):
      BIT-EXACT codes and scales - the north-star's first correctness gate;
  (b) the CPU oracle (oracle/sage_oracle.py) on the same seeded inputs.
"""
import pytest
import torch

import sageattention_b200 as sb
from oracle import sage_oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def make(b, h, n, d, dtype, layout, seed=0, bias=2.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(b, h, n, d, generator=g) + bias * torch.randn(1, h, 1, d, generator=g)
    x = x.to(dtype).to(DEV)
    if layout == "NHD":
        x = x.transpose(1, 2).contiguous()
    return x


def ref_per_warp(ref_fused, q, k, km, BLKQ, WARPQ, BLKK, layout):
    """The reference's per_warp_int8 (quant.py:106-181) re-stated over the reference's compiled launchers."""
    tl_ = 0 if layout == "NHD" else 1
    if layout == "HND":
        b, hq, nq, d = q.shape
        _, hk, nk, _ = k.shape
    else:
        b, nq, hq, d = q.shape
        _, nk, hk, _ = k.shape
    q8 = torch.empty(q.shape, dtype=torch.int8, device=q.device)
    k8 = torch.empty(k.shape, dtype=torch.int8, device=k.device)
    qs = torch.zeros((b, hq, ((nq + BLKQ - 1) // BLKQ) * (BLKQ // WARPQ)), device=q.device, dtype=torch.float32)
    ks = torch.zeros((b, hk, (nk + BLKK - 1) // BLKK), device=q.device, dtype=torch.float32)
    ref_fused.quant_per_warp_int8_cuda(q, q8, qs, BLKQ, WARPQ, tl_)
    if km is not None:
        ref_fused.quant_per_block_int8_fuse_sub_mean_cuda(k, km, k8, ks, BLKK, tl_)
    else:
        ref_fused.quant_per_block_int8_cuda(k, k8, ks, BLKK, tl_)
    return q8, qs, k8, ks


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("n,d", [(256, 128), (1000, 128), (333, 64), (64, 64), (17, 128)])
@pytest.mark.parametrize("warpq", [32, 16])
def test_per_warp_bit_exact_vs_reference(ref_fused, dtype, layout, n, d, warpq):
    q = make(2, 3, n, d, dtype, layout, 1)
    k = make(2, 3, n + 37, d, dtype, layout, 2)
    km = k.mean(dim=2 if layout == "HND" else 1)
    for use_km in (km, None):
        mine = sb.per_warp_int8(q, k, use_km, WARPQ=warpq, tensor_layout=layout)
        ref = ref_per_warp(ref_fused, q, k, use_km, 128, warpq, 64, layout)
        nw = (n + warpq - 1) // warpq
        assert torch.equal(mine[0], ref[0]), "q codes differ"
        assert torch.equal(mine[1][..., :nw], ref[1][..., :nw]), "q scales differ"
        assert torch.equal(mine[2], ref[2]), "k codes differ"
        assert torch.equal(mine[3], ref[3]), "k scales differ"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("n,d", [(512, 128), (777, 64)])
def test_per_block_bit_exact_vs_reference(ref_fused, dtype, n, d):
    q = make(1, 4, n, d, dtype, "HND", 3)
    k = make(1, 4, n, d, dtype, "HND", 4)
    km = k.mean(dim=2)
    mine = sb.per_block_int8(q, k, km, sm_scale=0.11)
    q8 = torch.empty_like(q, dtype=torch.int8)
    k8 = torch.empty_like(k, dtype=torch.int8)
    qs = torch.empty((1, 4, (n + 127) // 128), device=DEV, dtype=torch.float32)
    ks = torch.empty((1, 4, (n + 63) // 64), device=DEV, dtype=torch.float32)
    ref_fused.quant_per_block_int8_sm_scale_cuda(q, q8, qs, 0.11 * 1.44269504, 128, 1)
    ref_fused.quant_per_block_int8_fuse_sub_mean_cuda(k, km, k8, ks, 64, 1)
    assert torch.equal(mine[0], q8) and torch.equal(mine[1], qs)
    assert torch.equal(mine[2], k8) and torch.equal(mine[3], ks)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("n,d", [(256, 128), (1000, 128), (130, 64), (50, 64)])
@pytest.mark.parametrize("smooth,scale_max", [(False, 448.0), (True, 448.0), (False, 2.25)])
def test_per_channel_fp8_bit_exact_vs_reference(ref_fused, dtype, layout, n, d, smooth, scale_max):
    v = make(2, 2, n, d, dtype, layout, 5, bias=1.0)
    v8, vs, vm = sb.per_channel_fp8(v, tensor_layout=layout, scale_max=scale_max, smooth_v=smooth)
    # reference pipeline (quant.py:270-322) over the reference's compiled launchers
    tl_ = 0 if layout == "NHD" else 1
    npad = (n + 63) // 64 * 64
    shape = (2, 2, d, npad) if layout == "HND" else (2, d, 2, npad)
    vt = torch.zeros(shape, dtype=dtype, device=DEV)
    ref_fused.transpose_pad_permute_cuda(v, vt, tl_)
    r8 = torch.zeros(shape, dtype=torch.float8_e4m3fn, device=DEV)
    rs = torch.zeros((2, 2, d), dtype=torch.float32, device=DEV)
    rm = torch.empty((2, 2, d), dtype=torch.float32, device=DEV)
    if smooth:
        ref_fused.mean_scale_fuse_quant_cuda(vt, r8, rm, rs, n, scale_max, tl_)
    else:
        ref_fused.scale_fuse_quant_cuda(vt, r8, rs, n, scale_max, tl_)
    assert v8.shape == r8.shape
    assert torch.equal(vs, rs), "v_scale differs"
    if smooth:
        assert torch.equal(vm, rm), "v_mean differs"
    assert torch.equal(v8.view(torch.uint8), r8.view(torch.uint8)), "fp8 codes differ"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
def test_sub_mean_bit_exact_vs_reference(ref_fused, dtype, layout):
    v = make(2, 3, 300, 128, dtype, layout, 6)
    vm = v.mean(dim=2 if layout == "HND" else 1)
    mine, _ = sb.sub_mean(v, tensor_layout=layout, vm=vm)
    ref = torch.empty(v.shape, dtype=torch.float16, device=DEV)
    ref_fused.sub_mean_cuda(v, vm, ref, 0 if layout == "NHD" else 1)
    assert torch.equal(mine, ref)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("n,d", [(4096, 128), (1000, 64), (100, 128)])
def test_seq_mean_matches_torch(dtype, n, d):
    k = make(2, 4, n, d, dtype, "HND", 7)
    mine = sb.seq_mean(k)
    ref = k.mean(dim=2)
    # both reduce in fp32 and round once to the input dtype; the summation order differs, so a
    # value on a rounding boundary may land on the neighbouring code: <= 1 ulp, and rare
    diff = (mine.float() - ref.float()).abs()
    ulp = ref.float().abs() * (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7) + 1e-7
    assert (diff <= ulp).all()
    assert (mine != ref).float().mean().item() < 0.01


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("n,d,warpq", [(256, 128, 32), (1000, 128, 16), (333, 64, 32), (40, 64, 32)])
def test_quantizers_match_cpu_oracle(dtype, layout, n, d, warpq):
    """CUDA path vs CPU oracle.  The oracle divides in IEEE arithmetic, the CUDA reference (and
    therefore our kernels) with div.approx: scales may differ by 1 ulp and a code by +-1 when
    x*inv lands within an ulp of a rounding boundary.  Bound: <= 0.1% of codes, each by at most 1."""
    q = make(1, 2, n, d, dtype, layout, 8)
    k = make(1, 2, n + 5, d, dtype, layout, 9)
    km = k.mean(dim=2 if layout == "HND" else 1)
    for name in ("per_warp_int8", "per_thread_int8"):
        mine = getattr(sb, name)(q, k, km, WARPQ=warpq, tensor_layout=layout)
        orc = getattr(O, name)(q.cpu(), k.cpu(), km.cpu(), WARPQ=warpq, tensor_layout=layout)
        for i in (0, 2):
            dc = (mine[i].cpu().int() - orc[i].int()).abs()
            assert dc.max().item() <= 1, f"{name}[{i}] code off by more than 1"
            assert (dc != 0).float().mean().item() <= 1e-3, f"{name}[{i}] too many differing codes"
        for i in (1, 3):
            w = orc[i].shape[-1]
            a, bb = mine[i].cpu()[..., :w], orc[i]
            nz = bb != 0
            assert ((a[nz] - bb[nz]).abs() <= bb[nz].abs() * 2.4e-7).all(), f"{name}[{i}] scale beyond 2 ulp"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("gran", ["per_thread", "per_warp"])
@pytest.mark.parametrize("n,d", [(4096, 128), (1000, 64), (100, 128), (5003, 128), (20000, 64)])
def test_fused_k_smoothing_equals_separate_calls(dtype, layout, gran, n, d):
    """sage_smooth_quant_k_int8 (one cluster launch) must give the mean within 1 ulp of torch's and,
    for ITS mean, exactly the codes and scales of the stand-alone quantiser."""
    k = make(2, 3, n, d, dtype, layout, 11) + 0.75
    q = make(2, 3, 128, d, dtype, layout, 12)
    k_int8, k_scale, km = sb.smooth_quant_k_int8(k, gran, layout)
    ref = k.mean(dim=2 if layout == "HND" else 1)
    diff = (km.float() - ref.float()).abs()
    ulp = ref.float().abs() * (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7) + 1e-7
    assert (diff <= ulp).all()
    fn = sb.per_thread_int8 if gran == "per_thread" else sb.per_warp_int8
    _, _, k_ref, s_ref = fn(q, k, km, tensor_layout=layout)
    assert torch.equal(k_int8, k_ref)
    assert torch.equal(k_scale, s_ref)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("gran,warpq", [("per_thread", 32), ("per_thread", 16), ("per_warp", 32), ("per_warp", 16)])
@pytest.mark.parametrize("nq,nk,d,hq,hk", [(4096, 4096, 128, 4, 4), (1000, 1500, 64, 4, 2), (77, 100, 128, 2, 2),
                                           (130, 5003, 128, 6, 2)])
def test_single_launch_qk_prologue_equals_separate_calls(dtype, layout, gran, warpq, nq, nk, d, hq, hk):
    """sage_prepass_qk_int8 (Q tiles and K clusters in one grid) = quant_q_int8 + smooth_quant_k_int8, bit for bit
    (GQA head counts and different q / kv lengths included)."""
    q = make(2, hq, nq, d, dtype, layout, 21)
    k = make(2, hk, nk, d, dtype, layout, 22) + 0.5
    q8, qs, k8, ks, km = sb.prepass_qk_int8(q, k, gran, WARPQ=warpq, tensor_layout=layout)
    q8r, qsr = sb.quant_q_int8(q, gran, WARPQ=warpq, tensor_layout=layout)
    k8r, ksr, kmr = sb.smooth_quant_k_int8(k, gran, layout)
    assert torch.equal(km, kmr)
    assert torch.equal(k8, k8r) and torch.equal(ks, ksr)
    assert torch.equal(q8, q8r)
    nwritten = (nq + warpq - 1) // warpq if gran == "per_warp" else qs.size(-1)
    assert torch.equal(qs[..., :nwritten], qsr[..., :nwritten])


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("gran", ["per_thread", "per_warp"])
@pytest.mark.parametrize("scale_max", [448.0, 2.25])
@pytest.mark.parametrize("nq,nk,d,hq,hk", [(4096, 4096, 128, 4, 4), (1000, 1500, 64, 4, 2), (77, 100, 128, 2, 2),
                                           (130, 5003, 128, 6, 2)])
def test_single_launch_qkv_prologue_equals_separate_calls(dtype, layout, gran, scale_max, nq, nk, d, hq, hk):
    """sage_prepass_qkv_fp8 (K slices, V slices and Q tiles in one grid) = sage_prepass_qk_int8 +
    sage_per_channel_fp8(smooth_v = 0), bit for bit, including the padded tail of v_fp8 and the caller-side
    padding to 128 tokens (n_stat)."""
    q = make(2, hq, nq, d, dtype, layout, 31)
    k = make(2, hk, nk, d, dtype, layout, 32) + 0.5
    v = make(2, hk, nk, d, dtype, layout, 33)
    n_stat = (nk + 127) // 128 * 128
    q8, qs, k8, ks, km, v8, vs = sb.prepass_qkv_fp8(q, k, v, gran, WARPQ=32, tensor_layout=layout,
                                                    scale_max=scale_max, _n_stat=n_stat)
    q8r, qsr, k8r, ksr, kmr = sb.prepass_qk_int8(q, k, gran, WARPQ=32, tensor_layout=layout)
    v8r, vsr, _ = sb.per_channel_fp8(v, tensor_layout=layout, scale_max=scale_max, smooth_v=False, _n_stat=n_stat)
    assert torch.equal(km, kmr)
    assert torch.equal(k8, k8r) and torch.equal(ks, ksr)
    assert torch.equal(q8, q8r)
    nwritten = (nq + 31) // 32 if gran == "per_warp" else qs.size(-1)
    assert torch.equal(qs[..., :nwritten], qsr[..., :nwritten])
    assert torch.equal(vs, vsr)
    assert v8.shape == v8r.shape and torch.equal(v8.view(torch.uint8), v8r.view(torch.uint8))


@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("gran,warpq", [("per_thread", 32), ("per_warp", 16)])
@pytest.mark.parametrize("nq,nk,d,hq,hk", [(4096, 4096, 128, 4, 4), (1000, 1500, 64, 4, 2), (77, 100, 128, 2, 2),
                                           (130, 5003, 96, 6, 2)])
def test_single_launch_qk_prologue_with_v_f16_copy(layout, gran, warpq, nq, nk, d, hq, hk):
    """sage_prepass_qk_int8_v16: the Q/K prologue launch also writes v.to(float16) for a bf16 V
    (/root/reference/sageattention/core.py:633).  Codes, scales and mean bit-identical to sage_prepass_qk_int8, the
    copy bit-identical to torch's cast (values beyond the f16 range included)."""
    dtype = torch.bfloat16
    q = make(2, hq, nq, d, dtype, layout, 51)
    k = make(2, hk, nk, d, dtype, layout, 52) + 0.5
    v = make(2, hk, nk, d, dtype, layout, 53)
    v.view(-1)[::1013] = 1.0e5        # beyond f16: the cast gives inf, and so must the kernel
    v.view(-1)[5::2027] = -3.0e-6     # f16 subnormal range
    fold = d not in (64, 128)
    base = sb.prepass_qk_int8(q, k, gran, WARPQ=warpq, tensor_layout=layout, _pad_head_dim=fold)
    out = sb.prepass_qk_int8(q, k, gran, WARPQ=warpq, tensor_layout=layout, _pad_head_dim=fold, v=v)
    assert len(out) == 6
    nwritten = (nq + warpq - 1) // warpq if gran == "per_warp" else base[1].size(-1)
    assert torch.equal(out[0], base[0]) and torch.equal(out[2], base[2]) and torch.equal(out[3], base[3])
    assert torch.equal(out[1][..., :nwritten], base[1][..., :nwritten]) and torch.equal(out[4], base[4])
    ref = v.to(torch.float16)
    assert out[5].shape == v.shape and out[5].dtype == torch.float16
    assert torch.equal(out[5].view(torch.int16), ref.view(torch.int16))


def test_fp16_pv_operator_with_bf16_input_launches_no_torch_copy():
    """With bf16 q, k, v the f16-P.V operator is two launches of this library (prologue incl. the V copy, attention)
    and nothing from torch: no at::direct_copy_kernel in the profile (the eager v.to(float16))."""
    q = make(1, 4, 1024, 128, torch.bfloat16, "HND", 61)
    k = make(1, 4, 1024, 128, torch.bfloat16, "HND", 62)
    v = make(1, 4, 1024, 128, torch.bfloat16, "HND", 63)
    ref = sb.sageattn_qk_int8_pv_fp16_cuda(q, k, v)   # warm-up
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        out = sb.sageattn_qk_int8_pv_fp16_cuda(q, k, v)
        torch.cuda.synchronize()
    names = [e.key for e in prof.key_averages()]
    assert not any("direct_copy" in n or "elementwise" in n for n in names), names
    assert any("prepass_qk_kernel" in n for n in names) and any("sage_attn_fwd_kernel" in n for n in names), names
    assert torch.equal(out, ref)
