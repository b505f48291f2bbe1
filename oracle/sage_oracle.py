"""TEST INFRASTRUCTURE - CPU restatement (torch CPU tensors, fp32/int arithmetic) of the reference's
hot path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product (sageattention_b200/) never does.

Parity status: the reference ships NO tests, golden vectors or known-answer fixtures for this path
 and none of it runs on a CPU, so this oracle is pinned differentially:
tests/golden/*.pt holds outputs of the reference's OWN CUDA sources (oracle/_ref/*.so, built by
oracle/build_ref.py from /root/reference) run on a B200 by tests/golden/make_golden.py, and
tests/test_oracle_golden.py checks every function below against them.  Where the reference binary
uses approximate hardware instructions (div.approx / MUFU) the CPU restatement uses IEEE
arithmetic; the documented consequence is scales within 1 ulp and a bounded number of +-1 codes
(stated in the test), everything else bit-exact.

Each function cites the reference lines it follows (paths relative to /root/reference).
"""
import math

import torch

PERM16 = [0, 1, 8, 9, 2, 3, 10, 11, 4, 5, 12, 13, 6, 7, 14, 15]
LOG2E = 1.44269504
FP8_OFFSET = 8.807  # csrc/qattn/attn_utils.cuh:30


def _to_bhnd(x, tensor_layout):
    return x if tensor_layout == "HND" else x.transpose(1, 2)


def _from_bhnd(x, tensor_layout):
    return x if tensor_layout == "HND" else x.transpose(1, 2)


# --------------------------------------------------------------------------------------------
# INT8 quantizers
# --------------------------------------------------------------------------------------------
def quant_int8_block(x, block, km=None, sm_scale=None, tensor_layout="HND"):
    """QuantInt8Kernel, csrc/fused/fused.cu:64-198 (launchers :429-768).
    x' = float(x) [- float(km)] [* sm_scale]; amax = max(1e-7, max|x'|) per `block` rows;
    scale = amax / 127; code = cvt.rni.sat.s8(x' * (127 / amax)).  Rows past N count as zeros."""
    xb = _to_bhnd(x, tensor_layout).float()
    b, h, n, d = xb.shape
    if km is not None:
        xb = xb - km.reshape(b, h, 1, d).float()
    if sm_scale is not None:
        xb = xb * torch.tensor(sm_scale, dtype=torch.float32)
    nblk = (n + block - 1) // block
    pad = nblk * block - n
    xp = torch.nn.functional.pad(xb, (0, 0, 0, pad)) if pad else xb
    xp = xp.reshape(b, h, nblk, block, d)
    amax = xp.abs().amax(dim=(3, 4)).clamp_min(torch.tensor(1e-7, dtype=torch.float32))
    scale = amax / 127.0
    inv = (127.0 / amax)[..., None, None]
    codes = torch.round(xp * inv).clamp_(-128, 127).to(torch.int8).reshape(b, h, nblk * block, d)[:, :, :n]
    return _from_bhnd(codes, tensor_layout).contiguous(), scale.contiguous()


def per_block_int8(q, k, km=None, BLKQ=128, BLKK=64, sm_scale=None, tensor_layout="HND"):
    """sageattention/quant.py:23-104."""
    if sm_scale is None:
        sm_scale = q.size(-1) ** -0.5
    q8, qs = quant_int8_block(q, BLKQ, None, sm_scale * 1.44269504, tensor_layout)
    k8, ks = quant_int8_block(k, BLKK, km, None, tensor_layout)
    return q8, qs, k8, ks


def per_warp_int8(q, k, km=None, BLKQ=128, WARPQ=32, BLKK=64, tensor_layout="HND"):
    """sageattention/quant.py:106-181.  q_scale is returned with ceil(N/BLKQ)*(BLKQ/WARPQ) entries;
    entries past ceil(N/WARPQ) are unwritten in the reference (torch.empty) and 0 here."""
    q8, qs = quant_int8_block(q, WARPQ, None, None, tensor_layout)
    n = q.size(2) if tensor_layout == "HND" else q.size(1)
    full = ((n + BLKQ - 1) // BLKQ) * (BLKQ // WARPQ)
    if qs.size(-1) < full:
        qs = torch.nn.functional.pad(qs, (0, full - qs.size(-1)))
    k8, ks = quant_int8_block(k, BLKK, km, None, tensor_layout)
    return q8, qs, k8, ks


def _quant_groups_triton(xg):
    """xg: [..., rows_in_group, D] fp32 -> (codes, scale) with the Triton arithmetic of
    sageattention/triton/quant_per_thread.py:39-44: scale = max|x|/127 + 1e-7; y = x/scale;
    y += 0.5*sign; truncate."""
    amax = xg.abs().amax(dim=(-1, -2))
    # the reference binary evaluates `amax / 127. + 1e-7` as ONE fma with the f32 reciprocal of 127
    # (ptxas folds Triton's div.full.f32-by-constant + add.f32; see prepass.cu pt_scale).  The f64
    # product of two f32 values is exact, so rounding the f64 sum to f32 reproduces the fma.
    rcp127 = torch.tensor([0x3C010204], dtype=torch.int32).view(torch.float32).double()
    scale = (amax.double() * rcp127 + torch.tensor(1e-7, dtype=torch.float32).double()).float()
    y = xg / scale[..., None, None]
    y = y + torch.where(y >= 0, 0.5, -0.5)
    return torch.trunc(y).to(torch.int8), scale


def per_thread_int8(q, k, km=None, BLKQ=128, WARPQ=32, BLKK=64, WARPK=64, tensor_layout="HND"):
    """sageattention/triton/quant_per_thread.py:22-102,158-208.
    Q groups: rows {r, r+8, ...} of each WARPQ tile (8 scales per tile);
    K groups: rows {8i+2c, 8i+2c+1} of each 64-row tile (4 scales), after k <- dtype(k - km)."""
    qb = _to_bhnd(q, tensor_layout)
    kb = _to_bhnd(k, tensor_layout)
    b, hq, nq, d = qb.shape
    _, hk, nk, _ = kb.shape
    if km is not None:
        kb = kb - km.reshape(b, hk, 1, d)  # in the input dtype, :162-163
    # ---- Q
    ntile = (nq + BLKQ - 1) // BLKQ * (BLKQ // WARPQ)
    qp = torch.nn.functional.pad(qb.float(), (0, 0, 0, ntile * WARPQ - nq))
    qg = qp.reshape(b, hq, ntile, WARPQ // 8, 8, d).permute(0, 1, 2, 4, 3, 5)  # [.., tile, r, i, d]
    qc, qs = _quant_groups_triton(qg)
    q8 = qc.permute(0, 1, 2, 4, 3, 5).reshape(b, hq, ntile * WARPQ, d)[:, :, :nq]
    q_scale = qs.reshape(b, hq, ntile * 8)
    # ---- K
    ktile = (nk + BLKK - 1) // BLKK * (BLKK // WARPK)
    kp = torch.nn.functional.pad(kb.float(), (0, 0, 0, ktile * WARPK - nk))
    kg = kp.reshape(b, hk, ktile, WARPK // 8, 4, 2, d).permute(0, 1, 2, 4, 3, 5, 6)  # [.., tile, c, i, two, d]
    kg2 = kg.reshape(b, hk, ktile, 4, (WARPK // 8) * 2, d)
    kc, ks = _quant_groups_triton(kg2)
    k8 = kc.reshape(b, hk, ktile, 4, WARPK // 8, 2, d).permute(0, 1, 2, 4, 3, 5, 6).reshape(b, hk, ktile * WARPK, d)[:, :, :nk]
    k_scale = ks.reshape(b, hk, ktile * 4)
    return (_from_bhnd(q8, tensor_layout).contiguous(), q_scale.contiguous(),
            _from_bhnd(k8, tensor_layout).contiguous(), k_scale.contiguous())


# --------------------------------------------------------------------------------------------
# V path
# --------------------------------------------------------------------------------------------
def sub_mean(v, tensor_layout="HND", vm=None):
    """SubMeanKernel, csrc/fused/fused.cu:200-260 + sageattention/quant.py:183-223:
    vm = v.mean(seq) in the input dtype; out = half(v - vm) with the subtraction in the input dtype."""
    vb = _to_bhnd(v, tensor_layout)
    if vm is None:
        vm = vb.mean(dim=2)
    out = (vb - vm.unsqueeze(2)).to(torch.float16)
    return _from_bhnd(out, tensor_layout).contiguous(), vm


def transpose_pad_permute(v, tensor_layout="HND"):
    """TransposePadPermuteKernel, csrc/fused/fused.cu:262-313: [.,N,D] -> [.,D,Npad64], zero pad,
    out[16g + j] = in[16g + PERM16[j]]."""
    vb = _to_bhnd(v, tensor_layout)
    b, h, n, d = vb.shape
    npad = (n + 63) // 64 * 64
    x = torch.zeros((b, h, npad, d), dtype=v.dtype)
    x[:, :, :n] = vb
    x = x.reshape(b, h, npad // 16, 16, d)[:, :, :, PERM16].reshape(b, h, npad, d)
    return x.permute(0, 1, 3, 2).contiguous()  # [B,H,D,Npad]


def _f32_to_e4m3_satfinite(x):
    """cvt.rn.satfinite.e4m3x2.f32 (csrc/numeric_conversion.cuh:46-60): RN-even, clamp to +-448, NaN kept."""
    xc = x.clamp(-448.0, 448.0)
    return xc.to(torch.float8_e4m3fn)


def _block_sum_reference_order(rows):
    """Sum over the last dim of an fp32 tensor [..., n] in the order of MeanScaleKernel
    (fused.cu:338-372) + vllm::blockReduceSum (csrc/reduction_utils.cuh:49-70) with 256 threads:
    thread t adds, in order, elements [8(256 i + t), +8) for i = 0, 1, ...; xor-shuffle tree
    (16,8,4,2,1) inside each warp; the 8 warp sums go through a second tree."""
    lead = rows.shape[:-1]
    n = rows.shape[-1]
    npk = (n + 7) // 8
    iters = (npk + 255) // 256
    buf = torch.zeros(lead + (iters * 256 * 8,), dtype=torch.float32)
    buf[..., :n] = rows
    buf = buf.reshape(lead + (iters, 256, 8))
    acc = torch.zeros(lead + (256,), dtype=torch.float32)
    for i in range(iters):
        for j in range(8):
            acc = acc + buf[..., i, :, j]

    def tree(vals):  # vals [..., 32]
        for m in (16, 8, 4, 2, 1):
            idx = torch.arange(32) ^ m
            vals = vals + vals[..., idx]
        return vals

    w = tree(acc.reshape(lead + (8, 32)))[..., 0]
    second = torch.zeros(lead + (32,), dtype=torch.float32)
    second[..., :8] = w
    return tree(second)[..., 0]


def per_channel_fp8(v, tensor_layout="HND", scale_max=448.0, smooth_v=True, n_stat=None):
    """sageattention/quant.py:225-322 + MeanScaleKernel csrc/fused/fused.cu:316-427 (e4m3fn,
    upstream NVIDIA semantics).  `n_stat`: pretend v was zero-padded to that many tokens first
    (core.py:822-828)."""
    vb = _to_bhnd(v, tensor_layout)
    b, h, n, d = vb.shape
    if n_stat is not None and n_stat > n:
        vb = torch.nn.functional.pad(vb, (0, 0, 0, n_stat - n))
        n = n_stat
    vt = transpose_pad_permute(vb, "HND").float()  # [B,H,D,Npad]
    npad = vt.size(-1)
    n16 = (n + 15) // 16 * 16
    mx = vt[..., :n16].amax(dim=-1)
    mn = vt[..., :n16].amin(dim=-1)
    if smooth_v:
        mean = _block_sum_reference_order(vt[..., :n16]) / n16
        amax = torch.maximum((mx - mean).abs(), (mn - mean).abs())
        y = (vt - mean[..., None]) * (scale_max / amax)[..., None]
    else:
        mean = None
        amax = torch.maximum(mx.abs(), mn.abs())
        y = vt * (scale_max / amax)[..., None]
    v_scale = amax / scale_max
    v8 = _f32_to_e4m3_satfinite(y)
    if tensor_layout == "NHD":
        v8 = v8.permute(0, 2, 1, 3).contiguous()  # [B,D,H,Npad]
    return v8, v_scale, mean


# --------------------------------------------------------------------------------------------
# Attention kernels (tile-by-tile online softmax, the reference's order of operations)
# --------------------------------------------------------------------------------------------
def _scale_index(n_rows, gran, is_q, warp_q=32):
    """Row -> scale index.  per_warp (2): q row//WARPQ, k row//64 (qk_int_sv_f16_cuda_sm80.cu:100-110);
    per_thread (3): q (row//WARPQ)*8 + row%8, k (row//64)*4 + (row%8)//2 (:111-132,
    quant_per_thread.py:32,75-76); per_block (1): q row//128, k row//64."""
    r = torch.arange(n_rows)
    if is_q:
        if gran == 3:
            return (r // warp_q) * 8 + r % 8
        if gran == 2:
            return r // warp_q
        return r // 128
    if gran == 3:
        return (r // 64) * 4 + (r % 8) // 2
    return r // 64


def attn_int8(q8, k8, v, q_scale, k_scale, *, pv="f16", v_scale=None, v_mean=None, gran=3, warp_q=32,
              is_causal=False, sm_scale=None, return_lse=False, out_dtype=torch.float16, cta_k=64):
    """qk_int_sv_f16_attn_kernel (csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671) for pv="f16" and
    qk_int_sv_f8_attn_kernel (csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh:44-707) for pv="f8", HND tensors.

    Per 64-row KV tile (csrc/qattn/attn_utils.cuh:355-458 update_mdo):
      S = q8 . k8^T (exact s32) -> f32;  s' = S * (q_scale*k_scale) * (sm_scale*log2e)
      masked (kv > q if causal, kv >= kv_len): -5e6                       (:297-351)
      m_new = max(m, rowmax s')  [f8: the exponent is biased by +8.807, :377-380]
      O *= 2^(m - m_new); d *= 2^(m - m_new);  p = 2^(s' - m_new [+ 8.807])
      f16: P -> fp16 RN, d += rowsum(fp16 P)  (:461-476, :546)   O += P.V  (fp16 x fp16 -> f32)
      f8 : d += rowsum(f32 p)  (:548-553), P -> e4m3 RN-satfinite (:479-493), O += P.V8 (f32)
    out = O / d [* v_scale] [+ v_mean]  (:724-775, sm89.cuh:578-624);  lse = log2(d) + m [- 8.807]
    v: fp16 [B,Hkv,Nk,D] (f16) or e4m3 [B,Hkv,D,Npad] in the permuted layout (f8).
    m starts at -5e6 and d at 1 in the reference (sm80.cu:161-170); the first rescale zeroes d.
    """
    b, hq, nq, d = q8.shape
    _, hk, nk, _ = k8.shape
    g = hq // hk
    if sm_scale is None:
        sm_scale = d ** -0.5
    sml2 = torch.tensor(sm_scale * LOG2E, dtype=torch.float32)
    qi = _scale_index(nq, gran, True, warp_q).clamp_max(q_scale.size(-1) - 1)
    ki = _scale_index(nk, gran, False).clamp_max(k_scale.size(-1) - 1)
    qs_row = q_scale[:, :, qi]                                   # [B,Hq,Nq]
    ks_col = k_scale[:, :, ki].repeat_interleave(g, dim=1)      # [B,Hq,Nk]
    qf = q8.float()
    kf = k8.float().repeat_interleave(g, dim=1)
    if pv == "f16":
        vf = v.float().repeat_interleave(g, dim=1)              # [B,Hq,Nk,D]
    else:
        inv = torch.tensor(PERM16).argsort()
        npad = v.size(-1)
        vt = v.float().reshape(b, hk, d, npad // 16, 16)[..., inv].reshape(b, hk, d, npad)
        vf = vt.transpose(2, 3)[:, :, :nk].repeat_interleave(g, dim=1)   # un-permuted [B,Hq,Nk,D]
    off = FP8_OFFSET if pv == "f8" else 0.0
    m = torch.full((b, hq, nq), -5000000.0, dtype=torch.float32)
    dsum = torch.ones((b, hq, nq), dtype=torch.float32)
    o = torch.zeros((b, hq, nq, d), dtype=torch.float32)
    qidx = torch.arange(nq)[:, None]
    for k0 in range(0, nk, cta_k):
        k1 = min(nk, k0 + cta_k)
        s = torch.matmul(qf, kf[:, :, k0:k1].transpose(2, 3))             # exact in fp32 (|S| < 2^24)
        s = s * (qs_row[..., None] * ks_col[:, :, None, k0:k1])
        if is_causal:
            kvidx = torch.arange(k0, k1)[None, :]
            s = torch.where(kvidx > qidx, torch.tensor(-5000000.0), s)
        s = s * sml2
        m_new = torch.maximum(m, s.amax(dim=-1))
        alpha = torch.exp2(m - m_new)
        p = torch.exp2(s - m_new[..., None] + off)
        if pv == "f16":
            p = p.to(torch.float16).float()
            dsum = dsum * alpha + p.sum(-1)
        else:
            dsum = dsum * alpha + p.sum(-1)
            p = _f32_to_e4m3_satfinite(p).float()
        o = o * alpha[..., None] + torch.matmul(p, vf[:, :, k0:k1])
        m = m_new
    out = o / dsum[..., None]
    if v_scale is not None:
        out = out * v_scale.repeat_interleave(g, dim=1)[:, :, None, :]
    if v_mean is not None:
        out = out + v_mean.float().repeat_interleave(g, dim=1)[:, :, None, :]
    out = out.to(out_dtype)
    if return_lse:
        return out, torch.log2(dsum) + m - off
    return out


def sageattn_ref(q, k, v, *, pv="f8", tensor_layout="HND", is_causal=False, qk_quant_gran="per_thread",
                 sm_scale=None, pv_accum_dtype=None, smooth_k=True, smooth_v=False, return_lse=False):
    """End-to-end restatement of sageattn_qk_int8_pv_fp16_cuda (sageattention/core.py:480-653, pv="f16")
    and sageattn_qk_int8_pv_fp8_cuda (core.py:656-905, upstream semantics, pv="f8") on CPU tensors."""
    dtype = q.dtype
    d_og = q.size(-1)
    if d_og < 64:
        pad = 64 - d_og
    elif 64 < d_og < 128:
        pad = 128 - d_og
    else:
        pad = 0
    if pad:
        q, k, v = (torch.nn.functional.pad(t, (0, pad)) for t in (q, k, v))
    if sm_scale is None:
        sm_scale = d_og ** -0.5
    qh, kh, vh = (_to_bhnd(t, tensor_layout) for t in (q, k, v))
    km = None
    lse_corr = None
    if smooth_k:
        km = kh.mean(dim=2)                                                   # core.py:612 (input dtype)
        if return_lse:
            g = qh.size(1) // kh.size(1)
            lse_corr = torch.matmul(qh, km.repeat_interleave(g, dim=1).unsqueeze(-1)).squeeze(-1).float()
    if pv_accum_dtype is None:
        pv_accum_dtype = "fp32" if pv == "f16" else "fp32+fp16"
    warp_q = 16 if (pv == "f16" and q.size(-1) == 128 and pv_accum_dtype == "fp16+fp32") else 32
    if qk_quant_gran == "per_warp":
        q8, qs, k8, ks = per_warp_int8(qh, kh, km, WARPQ=warp_q)
        gran = 2
    else:
        q8, qs, k8, ks = per_thread_int8(qh, kh, km, WARPQ=warp_q)
        gran = 3
    if pv == "f16":
        vm = None
        if smooth_v and pv_accum_dtype == "fp16":
            v16, vm = sub_mean(vh)
        else:
            v16 = vh.to(torch.float16)
        res = attn_int8(q8, k8, v16, qs, ks, pv="f16", v_mean=vm, gran=gran, warp_q=warp_q, is_causal=is_causal,
                        sm_scale=sm_scale, return_lse=return_lse, out_dtype=dtype)
    else:
        if pv_accum_dtype in ("fp32+fp32", "fp32+fp16"):
            smooth_v = False
        scale_max = 2.25 if pv_accum_dtype == "fp32+fp16" else 448.0
        nk = kh.size(2)
        v8, vs, vm = per_channel_fp8(vh, scale_max=scale_max, smooth_v=smooth_v, n_stat=(nk + 127) // 128 * 128)
        res = attn_int8(q8, k8, v8, qs, ks, pv="f8", v_scale=vs, v_mean=vm, gran=gran, is_causal=is_causal,
                        sm_scale=sm_scale, return_lse=return_lse, out_dtype=dtype)
    if return_lse:
        o, lse = res
        lse = lse / LOG2E
        if lse_corr is not None:
            lse = lse + lse_corr * sm_scale
        return _from_bhnd(o, tensor_layout)[..., :d_og], lse
    return _from_bhnd(res, tensor_layout)[..., :d_og]


def sageattn_cpu_emulation(q, k, v, is_causal=False, sm_scale=None, smooth_k=True):
    """BASELINE.json config #1: torch CPU emulation of SageAttention-1 (per-block INT8 Q/K with K-mean
    smoothing, FP16 P.V) on fp32 inputs.  Follows sageattention/triton/quant_per_block.py:22-47 (Q block 128
    pre-scaled by sm_scale*log2e, K block 64, scale = amax/127, round-half-away) and
    sageattention/triton/attn_qk_int8_per_block.py:22-73 (base-2 online softmax, p -> fp16, fp16 V,
    fp32 accumulate), HND layout."""
    b, h, n, d = q.shape
    if sm_scale is None:
        sm_scale = d ** -0.5
    if smooth_k:
        k = k - k.mean(dim=2, keepdim=True)

    def qblock(x, blk, pre):
        nb = (x.size(2) + blk - 1) // blk
        xp = torch.nn.functional.pad(x.float() * pre, (0, 0, 0, nb * blk - x.size(2))).reshape(b, -1, nb, blk, d)
        scale = xp.abs().amax(dim=(3, 4)) / 127.0
        y = xp / scale[..., None, None]
        y = torch.trunc(y + torch.where(y >= 0, 0.5, -0.5))
        return y.reshape(b, -1, nb * blk, d)[:, :, :x.size(2)], scale

    q8, qs = qblock(q, 128, sm_scale * 1.44269504)
    k8, ks = qblock(k, 64, 1.0)
    v16 = v.to(torch.float16).float()
    nq, nk = q.size(2), k.size(2)
    qs_row = qs.repeat_interleave(128, dim=-1)[..., :nq]
    m = torch.full((b, h, nq), -float("inf"))
    l = torch.ones((b, h, nq))
    acc = torch.zeros((b, h, nq, d))
    qidx = torch.arange(nq)[:, None]
    for j, k0 in enumerate(range(0, nk, 64)):
        k1 = min(nk, k0 + 64)
        s = torch.matmul(q8, k8[:, :, k0:k1].transpose(2, 3)) * qs_row[..., None] * ks[:, :, j, None, None]
        if is_causal:
            s = torch.where(torch.arange(k0, k1)[None, :] > qidx, torch.tensor(-1.0e6), s)
        m_new = torch.maximum(m, s.amax(-1))
        p = torch.exp2(s - m_new[..., None])
        alpha = torch.exp2(m - m_new)
        l = l * alpha + p.sum(-1)
        acc = acc * alpha[..., None] + torch.matmul(p.to(torch.float16).float(), v16[:, :, k0:k1])
        m = m_new
    return acc / l[..., None]


def sageattn_varlen_ref(q, k, v, cu_seqlens_q, cu_seqlens_k, is_causal=False, sm_scale=None, smooth_k=True):
    """CPU restatement of sageattn_varlen (sageattention/core.py:363-477): q [Tq, Hq, D], k / v [Tk, Hk, D]
    packed along the token axis, cu_seqlens_* [batch + 1].

    core.py:443-453 zero-pads head_dim to 64 / 128 (sm_scale keeps the original head_dim, :467-468);
    :459-460 converts V to fp16; :462-464 smooths K with its mean over ALL tokens of the pack, in the input
    dtype; triton/quant_per_block_varlen.py:60 quantises per block with the blocks restarting at every
    sequence (Q blocks of 128 pre-scaled by sm_scale * log2 e, K blocks of 64); the Triton attention kernels
    (attn_qk_int8_block_varlen.py:123, attn_qk_int8_per_block_causal_varlen.py:138) run the base-2 online
    softmax per sequence with fp16 P and V and fp32 accumulation - i.e. `sageattn_cpu_emulation` per sequence.
    Hq == Hk only (the emulation has no GQA broadcast); returns fp32 [Tq, Hq, D]."""
    d = q.size(-1)
    if sm_scale is None:
        sm_scale = d ** -0.5
    d_pad = 64 if d <= 64 else 128
    if smooth_k:
        k = k - k.mean(dim=0, keepdim=True)
    outs = []
    for i in range(len(cu_seqlens_q) - 1):
        q0, q1 = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        k0, k1 = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        if q1 == q0:
            continue
        hnd = lambda t, a, b: torch.nn.functional.pad(t[a:b].float(), (0, d_pad - d)).transpose(0, 1).unsqueeze(0)
        o = sageattn_cpu_emulation(hnd(q, q0, q1), hnd(k, k0, k1), hnd(v, k0, k1), is_causal=is_causal,
                                   sm_scale=sm_scale, smooth_k=False)
        outs.append(o[0, :, :, :d].transpose(0, 1))
    return torch.cat(outs)
