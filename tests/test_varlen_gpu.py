"""GPU parity of sageattn_varlen (reference core.py:363-477) on packed variable-length
batches against
  (a) the CPU oracle's restatement of the reference's per-block scheme (oracle/sage_oracle.py
      sageattn_cpu_emulation: Q blocks of 128 pre-scaled by sm_scale*log2e, K blocks of 64, fp16 P and V, fp32
      accumulation) run per sequence on K smoothed with the pack-wide mean, as the reference does,
  (b) fp32 scaled_dot_product_attention per sequence.
Tolerances (floating point): cos >= 0.9999 / rel-L1 <= 1e-2 against the oracle (the CUDA quantiser rounds
half-to-even like fused.cu, the Triton one the reference uses here rounds half away: codes differ by 1 on
exact ties only); cos >= 0.999 and rel-L1 no worse than 1.15x the oracle's own against fp32 attention."""
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


@pytest.mark.parametrize("causal", [False, True])
def test_varlen_equal_lengths_batched_equals_per_sequence(causal):
    """A pack whose sequences all have the same lengths runs as ONE [B, n, H, D] batch (3 launches instead of
    3 per sequence); the result must be bit-identical to the per-sequence launches."""
    lens = [333] * 5
    q, k, v, cq, ck = _pack(lens, lens, 4, 2, 128, torch.bfloat16, seed=21)
    args = (q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV), 333, 333)
    o_batched = sb.sageattn_varlen(*args, is_causal=causal)
    o_loop = sb.sageattn_varlen(*args, is_causal=causal, _batched=False)
    torch.cuda.synchronize()
    assert torch.equal(o_batched, o_loop)
