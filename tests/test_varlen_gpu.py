"""GPU parity of sageattn_varlen (reference core.py:363-477) on packed variable-length
batches against
  (a) the CPU oracle's restatement of the reference's per-block scheme (oracle/sage_oracle.py
      sageattn_cpu_emulation: Q blocks of 128 pre-scaled by sm_scale*log2e, K blocks of 64, fp16 P and V, fp32
      accumulation) run per sequence on K smoothed with the pack-wide mean, as the reference does,
  (b) fp32 scaled_dot_product_attention per sequence.
  (c) the reference's OWN Triton code run on the same GPU (oracle/_ref/stage): its varlen quantiser
      (triton/quant_per_block_varlen.py) - codes, scales and prefix arrays BIT-EXACT - and its varlen attention
      kernels (attn_qk_int8_block_varlen.py, attn_qk_int8_per_block_causal_varlen.py) on the same codes.
Tolerances (floating point): cos >= 0.9999 / rel-L1 <= 1e-2 against the oracle and the reference kernels;
cos >= 0.999 and rel-L1 no worse than 1.15x the oracle's own against fp32 attention."""
import pytest
import torch

import sageattention_b200 as sb
from oracle import sage_oracle as O
from conftest import accuracy

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _pack(lens_q, lens_k, hq, hk, d, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, hq, d, generator=g)
    k = torch.randn(tk, hk, d, generator=g) + 2.0 * torch.randn(1, hk, d, generator=g)
    v = torch.randn(tk, hk, d, generator=g)
    cq = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    ck = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    return q.to(dtype), k.to(dtype), v.to(dtype), cq, ck


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("d", [128, 64, 96])
@pytest.mark.parametrize("causal", [False, True])
def test_varlen_vs_oracle_and_sdpa(dtype, d, causal):
    lens = [300, 128, 1000, 77, 1]
    h = 3
    q, k, v, cq, ck = _pack(lens, lens, h, h, d, dtype, seed=d + int(causal))
    o = sb.sageattn_varlen(q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV), max(lens), max(lens),
                           is_causal=causal)
    torch.cuda.synchronize()
    assert o.shape == q.shape and o.dtype == dtype
    oracle = O.sageattn_varlen_ref(q, k, v, cq, ck, is_causal=causal)
    refs = []
    for i in range(len(lens)):
        s0, s1 = int(cq[i]), int(cq[i + 1])
        f = lambda t: t[s0:s1].float().transpose(0, 1).unsqueeze(0)
        refs.append(torch.nn.functional.scaled_dot_product_attention(f(q), f(k), f(v), is_causal=causal)[0].transpose(0, 1))
    truth = torch.cat(refs)
    cos, l1, _ = accuracy(o.cpu(), oracle)
    assert cos >= 0.9999 and l1 <= 1e-2, (cos, l1)
    cos_t, l1_t, _ = accuracy(o.cpu(), truth)
    _, l1_o, _ = accuracy(oracle, truth)
    assert cos_t >= 0.999 and l1_t <= max(1.15 * l1_o, 1e-2), (cos_t, l1_t, l1_o)


def test_varlen_gqa_cross_lengths_and_errors():
    lens_q, lens_k = [200, 0, 513], [333, 64, 700]      # an empty query sequence; kv lengths differ from q lengths
    hq, hk, d = 8, 2, 128
    q, k, v, cq, ck = _pack(lens_q, lens_k, hq, hk, d, torch.bfloat16, seed=9)
    o = sb.sageattn_varlen(q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k))
    torch.cuda.synchronize()
    refs = []
    for i in range(len(lens_q)):
        a0, a1, b0, b1 = int(cq[i]), int(cq[i + 1]), int(ck[i]), int(ck[i + 1])
        if a1 == a0:
            continue
        f = lambda t, s0, s1, rep: t[s0:s1].float().transpose(0, 1).repeat_interleave(rep, 0).unsqueeze(0)
        refs.append(torch.nn.functional.scaled_dot_product_attention(
            f(q, a0, a1, 1), f(k, b0, b1, hq // hk), f(v, b0, b1, hq // hk))[0].transpose(0, 1))
    cos, l1, _ = accuracy(o.cpu(), torch.cat(refs))
    assert cos >= 0.999 and l1 <= 2e-2, (cos, l1)
    with pytest.raises(AssertionError):
        sb.sageattn_varlen(q, k, v, cq, ck, 1, 1)            # CPU tensors
    with pytest.raises(AssertionError):
        sb.sageattn_varlen(q.to(DEV).float(), k.to(DEV).float(), v.to(DEV).float(), cq.to(DEV), ck.to(DEV), 1000, 1000)


def _dev_pack(lens_q, lens_k, hq, hk, d, dtype, seed):
    q, k, v, cq, ck = _pack(lens_q, lens_k, hq, hk, d, dtype, seed)
    return q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("d", [128, 64])
@pytest.mark.parametrize("lens_q,lens_k,hq,hk", [
    ([300, 128, 1000, 77, 1], [300, 128, 1000, 77, 1], 3, 3),
    ([200, 0, 513, 256], [333, 64, 700, 64], 8, 2),          # empty query sequence, GQA, q / kv lengths differ
    ([4096], [4096], 2, 2),
    ([1, 1, 1, 129, 65], [1, 63, 64, 65, 127], 2, 1),
])
def test_varlen_quantizer_bit_exact_vs_reference_triton(ref_triton, dtype, d, lens_q, lens_k, hq, hk):
    """sage_quant_per_block_varlen_int8 against triton/quant_per_block_varlen.py:59-107 on the same inputs: codes,
    scales (the rows the reference has) and cu_seqlens_*_scale must be identical, with and without the folded
    `k - km` (the reference subtracts in torch first, core.py:462-463)."""
    tri = ref_triton("quant_per_block_varlen")
    q, k, v, cq, ck = _dev_pack(lens_q, lens_k, hq, hk, d, dtype, seed=len(lens_q) + d)
    q[5 % q.size(0)] = 0                                   # an all-zero row
    mq, mk = max(lens_q), max(lens_k)
    sm = d ** -0.5
    km = k.mean(dim=0, keepdim=True)
    for use_km in (None, km):
        mine = sb.per_block_int8_varlen(q, k, cq, ck, mq, mk, sm_scale=sm, km=use_km)
        ref = tri.per_block_int8(q, k if use_km is None else k - use_km, cq, ck, mq, mk, sm_scale=sm)
        torch.cuda.synchronize()
        nq_rows, nk_rows = int(ref[4][-1]), int(ref[5][-1])
        assert torch.equal(mine[4].long(), ref[4].long()) and torch.equal(mine[5].long(), ref[5].long())
        assert torch.equal(mine[0], ref[0]), "q codes differ"
        assert torch.equal(mine[2], ref[2]), "k codes differ"
        assert ref[1].shape == (nq_rows, hq) and torch.equal(mine[1][:nq_rows], ref[1]), "q scales differ"
        assert torch.equal(mine[3][:nk_rows], ref[3]), "k scales differ"
    # a block of zeros: scale 0, codes 0 (0 * inf = NaN truncates to 0 in the reference's sequence too)
    z = torch.zeros_like(q)
    mine = sb.per_block_int8_varlen(z, k, cq, ck, mq, mk, sm_scale=sm)
    ref = tri.per_block_int8(z, k, cq, ck, mq, mk, sm_scale=sm)
    assert torch.equal(mine[0], ref[0]) and torch.equal(mine[1][:int(ref[4][-1])], ref[1])


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [128, 64])
@pytest.mark.parametrize("lens_q,lens_k,hq,hk", [([300, 128, 1000, 77, 1], [300, 128, 1000, 77, 1], 4, 4),
                                                 ([200, 513, 256], [333, 700, 256], 4, 2)])
def test_varlen_attention_vs_reference_triton_kernels(ref_triton, causal, d, lens_q, lens_k, hq, hk):
    """The single-launch varlen attention against the reference's Triton varlen kernels on the SAME codes and scales
    (attn_qk_int8_block_varlen.py:123 / attn_qk_int8_per_block_causal_varlen.py:138)."""
    if causal and lens_q != lens_k:
        pytest.skip("the reference's causal kernel assumes qo_len == kv_len per sequence")
    att = ref_triton("attn_qk_int8_per_block_causal_varlen" if causal else "attn_qk_int8_block_varlen")
    dtype = torch.float16
    q, k, v, cq, ck = _dev_pack(lens_q, lens_k, hq, hk, d, dtype, seed=17 + d)
    mq, mk = max(lens_q), max(lens_k)
    km = k.mean(dim=0, keepdim=True)
    q8, qs, k8, ks, cqs, cks, _ = sb.per_block_int8_varlen(q, k, cq, ck, mq, mk, km=km)
    nq_rows, nk_rows = int(cqs[-1]), int(cks[-1])
    ro = att.forward(q8, k8, v, cq, ck, mq, qs[:nq_rows].contiguous(), ks[:nk_rows].contiguous(), cqs, cks,
                     output_dtype=dtype)
    o = sb.sageattn_varlen(q, k, v, cq, ck, mq, mk, is_causal=causal)
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all()
    cos, l1, _ = accuracy(o, ro)
    assert cos >= 0.9999 and l1 <= 1e-2, (cos, l1)


def test_varlen_is_three_launches_no_host_sync_and_graph_capturable():
    """A 5-sequence ragged pack: the pack-wide K mean, ONE quantiser launch, ONE attention launch, none
    of them per sequence; nothing on the host depends on the lengths, so the call can be captured in a CUDA graph and
    replayed with DIFFERENT cu_seqlens contents in the same buffers."""
    lens = [300, 128, 1000, 77, 1]
    q, k, v, cq, ck = _dev_pack(lens, lens, 4, 2, 128, torch.float16, seed=31)
    ref = sb.sageattn_varlen(q, k, v, cq, ck, 1000, 1000)
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        sb.sageattn_varlen(q, k, v, cq, ck, 1000, 1000)
        torch.cuda.synchronize()
    kernels = [(e.key, e.count) for e in prof.key_averages() if "Memset" not in e.key and "Memcpy" not in e.key]
    ours = [(n, c) for n, c in kernels if any(t in n for t in ("col_stats", "varlen_kernel", "attn_fwd"))]
    launched = [c for n, c in kernels if n == "cudaLaunchKernel"]
    assert sorted(c for _, c in ours) == [1, 1, 1], kernels
    assert launched in ([], [3]), kernels        # three kernel launches in all
    assert any("col_stats" in n for n, _ in ours) and any("varlen" in n for n, _ in ours) and \
        any("attn_fwd" in n for n, _ in ours), kernels
    # capture once, replay on a re-packed batch of the same total size
    o = torch.empty_like(q)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        sb.sageattn_varlen(q, k, v, cq, ck, 1000, 1000)
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        o_g = sb.sageattn_varlen(q, k, v, cq, ck, 1000, 1000)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(o_g, ref)
    lens2 = [1000, 306, 100, 100, 0]
    cq2 = torch.tensor([0] + list(torch.tensor(lens2).cumsum(0)), dtype=torch.int32, device=DEV)
    ref2 = sb.sageattn_varlen(q, k, v, cq2, cq2.clone(), 1000, 1000)
    cq.copy_(cq2)
    ck.copy_(cq2)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(o_g, ref2)


def test_varlen_sequence_without_keys_gives_zeros():
    lens_q, lens_k = [130, 200, 64], [256, 0, 100]
    q, k, v, cq, ck = _dev_pack(lens_q, lens_k, 2, 2, 64, torch.bfloat16, seed=41)
    o = sb.sageattn_varlen(q, k, v, cq, ck, 200, 256)
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all()
    assert (o[130:330] == 0).all()
    f = lambda t, a, b_: t[a:b_].float().transpose(0, 1).unsqueeze(0)
    for (a0, a1, b0, b1) in ((0, 130, 0, 256), (330, 394, 256, 356)):
        ref = torch.nn.functional.scaled_dot_product_attention(f(q, a0, a1), f(k, b0, b1), f(v, b0, b1))[0].transpose(0, 1)
        cos, l1, _ = accuracy(o[a0:a1], ref)
        assert cos >= 0.999 and l1 <= 2e-2, (cos, l1)
