"""Quantizers with the reference's signatures, backed by the sm_100a C-ABI library.

Mirrors /root/reference/sageattention/quant.py (per_block_int8 :23, per_warp_int8 :106,
sub_mean :183, per_channel_fp8 :225) and the Triton default-granularity quantizer
/root/reference/sageattention/triton/quant_per_thread.py:158 (per_thread_int8), which the
reference's CUDA entry points use by default and which exists here as a CUDA kernel because this
package has no Triton path.

Outputs are bit-exact with the reference quantizers on identical inputs (tests/test_quant_gpu.py
checks this against the reference's own compiled sources).  No CPU fallback exists.
"""
from typing import Optional

import torch

from . import _lib


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float16:
        return _lib.DTYPE_F16
    if t.dtype == torch.bfloat16:
        return _lib.DTYPE_BF16
    raise ValueError(f"Unsupported dtype {t.dtype}: tensors must be torch.float16 or torch.bfloat16")


def _dims(t: torch.Tensor, tensor_layout: str):
    """-> (B, H, N, D, Tensor4) for a [B,H,N,D] ("HND") or [B,N,H,D] ("NHD") tensor."""
    if t.dim() != 4:
        raise RuntimeError("Tensor must have 4 dimensions")
    if t.stride(-1) != 1:
        raise RuntimeError("Tensor must be contiguous at last dimension")
    if tensor_layout == "HND":
        b, h, n, d = t.shape
        t4 = _lib.Tensor4(t.data_ptr(), t.stride(0), t.stride(1), t.stride(2))
    elif tensor_layout == "NHD":
        b, n, h, d = t.shape
        t4 = _lib.Tensor4(t.data_ptr(), t.stride(0), t.stride(2), t.stride(1))
    else:
        raise ValueError(f"Unknown tensor layout: {tensor_layout}")
    return b, h, n, d, t4


def _stream(t: torch.Tensor):
    # raw handle of torch's current stream on the tensor's device (the Stream object costs ~5 us to build)
    return torch._C._cuda_getCurrentRawStream(t.device.index)


class _on_device:
    """`with torch.cuda.device(dev)` that costs nothing when `dev` already is the current device (the usual
    case; the context manager itself is ~5 us per call)."""
    __slots__ = ("idx", "prev")

    def __init__(self, dev: torch.device):
        self.idx = dev.index
        self.prev = -1

    def __enter__(self):
        cur = torch._C._cuda_getDevice()
        if cur != self.idx:
            self.prev = cur
            torch._C._cuda_setDevice(self.idx)

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch._C._cuda_setDevice(self.prev)
        return False


def _require_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("Tensor must be on CUDA (sageattention_b200 has no CPU path)")


def _mean_ptr(km: Optional[torch.Tensor], b, h, d, like: torch.Tensor):
    if km is None:
        return None, None
    km = km.reshape(b, h, d)
    if km.dtype != like.dtype:
        raise RuntimeError("km must have the same dtype as the tensor it smooths")
    km = km.contiguous()
    return km, km.data_ptr()


def seq_mean(x: torch.Tensor, tensor_layout: str = "HND") -> torch.Tensor:
    """`x.mean(dim=seq)` in one deterministic kernel -> [B, H, D], same dtype as x
    (replaces the torch reduction at /root/reference/sageattention/core.py:612)."""
    _require_cuda(x)
    lib = _lib.load()
    b, h, n, d, x4 = _dims(x, tensor_layout)
    out = torch.empty((b, h, d), dtype=x.dtype, device=x.device)
    ws = torch.empty(lib.sage_seq_mean_workspace_bytes(b, h, n, d), dtype=torch.uint8, device=x.device)
    with _on_device(x.device):
        _lib.check(lib.sage_seq_mean(x4, out.data_ptr(), b, h, n, d, _dtype_code(x), ws.data_ptr(), _stream(x)))
    return out


def _quant_block(x, km, block, sm_scale, tensor_layout, scale_per_bh=None):
    lib = _lib.load()
    b, h, n, d, x4 = _dims(x, tensor_layout)
    out = torch.empty(x.shape, dtype=torch.int8, device=x.device)
    _, _, _, _, o4 = _dims(out, tensor_layout)
    nblk = (n + block - 1) // block
    if scale_per_bh is None:
        scale_per_bh = nblk
    # entries past ceil(N/block) are never written by the reference either (torch.empty there);
    # zeros keep them harmless
    scale = (torch.empty if scale_per_bh == nblk else torch.zeros)(
        (b, h, scale_per_bh), dtype=torch.float32, device=x.device)
    km_c, km_ptr = _mean_ptr(km, b, h, d, x)
    with _on_device(x.device):
        _lib.check(lib.sage_quant_int8(x4, km_ptr, o4, scale.data_ptr(), b, h, n, d, _dtype_code(x), block,
                                       scale_per_bh, 0 if sm_scale is None else 1,
                                       0.0 if sm_scale is None else float(sm_scale), _stream(x)))
    return out, scale


def per_block_int8(q, k, km=None, BLKQ=128, BLKK=64, sm_scale=None, tensor_layout="HND"):
    """Reference: /root/reference/sageattention/quant.py:23-104.  Q is pre-multiplied by
    sm_scale * 1.44269504 before quantisation; K is optionally smoothed by km."""
    _require_cuda(q, k, km)
    head_dim = q.size(-1)
    if sm_scale is None:
        sm_scale = head_dim ** -0.5
    sm_scale = sm_scale * 1.44269504
    q_int8, q_scale = _quant_block(q, None, BLKQ, sm_scale, tensor_layout)
    k_int8, k_scale = _quant_block(k, km, BLKK, None, tensor_layout)
    return q_int8, q_scale, k_int8, k_scale


def per_warp_int8(q, k, km=None, BLKQ=128, WARPQ=32, BLKK=64, tensor_layout="HND"):
    """Reference: /root/reference/sageattention/quant.py:106-181.  q_scale has
    ceil(Nq/BLKQ)*(BLKQ/WARPQ) entries per head (only the first ceil(Nq/WARPQ) are written,
    as in the reference, whose scale buffer is torch.empty)."""
    _require_cuda(q, k, km)
    n = q.size(2) if tensor_layout == "HND" else q.size(1)
    q_int8, q_scale = _quant_block(q, None, WARPQ, None, tensor_layout,
                                   scale_per_bh=((n + BLKQ - 1) // BLKQ) * (BLKQ // WARPQ))
    k_int8, k_scale = _quant_block(k, km, BLKK, None, tensor_layout)
    return q_int8, q_scale, k_int8, k_scale


def per_thread_int8(q, k, km=None, BLKQ=128, WARPQ=32, BLKK=64, WARPK=64, sm_scale=None, tensor_layout="HND"):
    """Reference: /root/reference/sageattention/triton/quant_per_thread.py:158-208 (Triton).
    Same group structure and arithmetic, as a CUDA kernel."""
    _require_cuda(q, k, km)
    if WARPK != 64 or BLKK != 64 or BLKQ != 128:
        raise ValueError("per_thread_int8 supports BLKQ=128, BLKK=64, WARPK=64 (the reference's only use)")
    lib = _lib.load()
    b, hq, nq, d, q4 = _dims(q, tensor_layout)
    _, hk, nk, _, k4 = _dims(k, tensor_layout)
    q_int8 = torch.empty(q.shape, dtype=torch.int8, device=q.device)
    k_int8 = torch.empty(k.shape, dtype=torch.int8, device=k.device)
    _, _, _, _, qo4 = _dims(q_int8, tensor_layout)
    _, _, _, _, ko4 = _dims(k_int8, tensor_layout)
    q_scale = torch.empty((b, hq, (nq + BLKQ - 1) // BLKQ * (BLKQ // WARPQ) * 8), dtype=torch.float32, device=q.device)
    k_scale = torch.empty((b, hk, (nk + BLKK - 1) // BLKK * (BLKK // WARPK) * 4), dtype=torch.float32, device=q.device)
    km_c, km_ptr = _mean_ptr(km, b, hk, d, k)
    with _on_device(q.device):
        _lib.check(lib.sage_quant_per_thread_int8(q4, None, qo4, q_scale.data_ptr(), b, hq, nq, d, _dtype_code(q),
                                                  0, WARPQ, _stream(q)))
        _lib.check(lib.sage_quant_per_thread_int8(k4, km_ptr, ko4, k_scale.data_ptr(), b, hk, nk, d, _dtype_code(k),
                                                  1, 64, _stream(k)))
    return q_int8, q_scale, k_int8, k_scale


def smooth_quant_k_int8(k, qk_quant_gran="per_thread", tensor_layout="HND"):
    """km = k.mean(dim=seq) and the K half of per_warp_int8 / per_thread_int8(…, km) in one launch
    (the prologue of /root/reference/sageattention/core.py:611-632 for K).  Returns
    (k_int8, k_scale, km[B,Hkv,D]); codes and scales are those the separate calls give for this km."""
    _require_cuda(k)
    lib = _lib.load()
    b, h, n, d, k4 = _dims(k, tensor_layout)
    k_int8 = torch.empty(k.shape, dtype=torch.int8, device=k.device)
    _, _, _, _, ko4 = _dims(k_int8, tensor_layout)
    per_thread = 1 if qk_quant_gran == "per_thread" else 0
    per_bh = (n + 63) // 64 * (4 if per_thread else 1)
    k_scale = torch.empty((b, h, per_bh), dtype=torch.float32, device=k.device)
    km = torch.empty((b, h, d), dtype=k.dtype, device=k.device)
    with _on_device(k.device):
        _lib.check(lib.sage_smooth_quant_k_int8(k4, km.data_ptr(), ko4, k_scale.data_ptr(), b, h, n, d,
                                                _dtype_code(k), per_thread, per_bh, _stream(k)))
    return k_int8, k_scale, km


def _padded_head_dim(d: int) -> int:
    """64 / 128 as core.py:590-601 pads to; raises for head_dim > 128."""
    if d > 128:
        raise ValueError(f"Unsupported head_dim: {d}")
    return 64 if d <= 64 else 128


def prepass_qk_int8(q, k, qk_quant_gran="per_thread", WARPQ=32, tensor_layout="HND", _pad_head_dim=False, v=None):
    """The operator's Q/K prologue with smooth_k in ONE launch (core.py:611-632): km = k.mean(seq),
    K codes/scales for that mean and Q codes/scales.  Returns (q_int8, q_scale, k_int8, k_scale, km),
    identical to quant_q_int8(q) + smooth_quant_k_int8(k).
    With `v` (same heads / length as k): the same launch also writes v16 = v.to(float16) (core.py:633) and the
    return value gains it as a sixth element."""
    _require_cuda(q, k)
    lib = _lib.load()
    b, hq, nq, d_og, q4 = _dims(q, tensor_layout)
    _, hk, nk, _, k4 = _dims(k, tensor_layout)
    # `_pad_head_dim` (internal): head_dim is a multiple of 8 but not 64 / 128 - the kernel reads the rows as
    # they are and writes codes / mean of the padded width (zeros beyond head_dim), instead of three F.pad copies
    d = _padded_head_dim(d_og) if _pad_head_dim else d_og
    per_thread = 1 if qk_quant_gran == "per_thread" else 0
    q_int8 = torch.empty(q.shape[:-1] + (d,), dtype=torch.int8, device=q.device)
    k_int8 = torch.empty(k.shape[:-1] + (d,), dtype=torch.int8, device=k.device)
    _, _, _, _, qo4 = _dims(q_int8, tensor_layout)
    _, _, _, _, ko4 = _dims(k_int8, tensor_layout)
    q_per_bh = (nq + 127) // 128 * (128 // WARPQ) * (8 if per_thread else 1)
    k_per_bh = (nk + 63) // 64 * (4 if per_thread else 1)
    # per_warp: entries past ceil(Nq/WARPQ) are never written by the reference either (torch.empty there)
    q_scale = (torch.empty if per_thread or q_per_bh == (nq + WARPQ - 1) // WARPQ else torch.zeros)(
        (b, hq, q_per_bh), dtype=torch.float32, device=q.device)
    k_scale = torch.empty((b, hk, k_per_bh), dtype=torch.float32, device=q.device)
    km = torch.empty((b, hk, d), dtype=k.dtype, device=k.device)
    if v is not None:
        _require_cuda(v)
        _, hv, nv, dv, v4 = _dims(v, tensor_layout)
        if (hv, nv, dv) != (hk, nk, d_og) or v.dtype != k.dtype:
            raise ValueError("k and v must have the same heads, length, head_dim and dtype")
        v16 = torch.empty(v.shape, dtype=torch.float16, device=v.device)
        _, _, _, _, vo4 = _dims(v16, tensor_layout)
        with _on_device(q.device):
            _lib.check(lib.sage_prepass_qk_int8_v16(q4, k4, qo4, ko4, q_scale.data_ptr(), k_scale.data_ptr(),
                                                    km.data_ptr(), v4, vo4, b, hq, hk, nq, nk, d, _dtype_code(q),
                                                    per_thread, WARPQ, q_per_bh, k_per_bh, d_og, _stream(q)))
        return q_int8, q_scale, k_int8, k_scale, km, v16
    with _on_device(q.device):
        _lib.check(lib.sage_prepass_qk_int8(q4, k4, qo4, ko4, q_scale.data_ptr(), k_scale.data_ptr(), km.data_ptr(),
                                            b, hq, hk, nq, nk, d, _dtype_code(q), per_thread, WARPQ, q_per_bh,
                                            k_per_bh, d_og, _stream(q)))
    return q_int8, q_scale, k_int8, k_scale, km


def prepass_qkv_fp8(q, k, v, qk_quant_gran="per_thread", WARPQ=32, tensor_layout="HND", scale_max=448.0,
                    _n_stat=None, _pad_head_dim=False):
    """The fp8-P.V operator's whole prologue (smooth_k, no smooth_v) in ONE launch.  Returns
    (q_int8, q_scale, k_int8, k_scale, km, v_fp8, v_scale), identical to prepass_qk_int8(q, k) +
    per_channel_fp8(v, smooth_v=False): K slices, V slices and Q tiles run side by side in one grid."""
    _require_cuda(q, k, v)
    lib = _lib.load()
    b, hq, nq, d_og, q4 = _dims(q, tensor_layout)
    _, hk, nk, _, k4 = _dims(k, tensor_layout)
    _, hv, nv, _, v4 = _dims(v, tensor_layout)
    if (hv, nv) != (hk, nk):
        raise ValueError("k and v must have the same heads and length")
    d = _padded_head_dim(d_og) if _pad_head_dim else d_og   # see prepass_qk_int8
    per_thread = 1 if qk_quant_gran == "per_thread" else 0
    q_int8 = torch.empty(q.shape[:-1] + (d,), dtype=torch.int8, device=q.device)
    k_int8 = torch.empty(k.shape[:-1] + (d,), dtype=torch.int8, device=k.device)
    _, _, _, _, qo4 = _dims(q_int8, tensor_layout)
    _, _, _, _, ko4 = _dims(k_int8, tensor_layout)
    q_per_bh = (nq + 127) // 128 * (128 // WARPQ) * (8 if per_thread else 1)
    k_per_bh = (nk + 63) // 64 * (4 if per_thread else 1)
    q_scale = (torch.empty if per_thread or q_per_bh == (nq + WARPQ - 1) // WARPQ else torch.zeros)(
        (b, hq, q_per_bh), dtype=torch.float32, device=q.device)
    k_scale = torch.empty((b, hk, k_per_bh), dtype=torch.float32, device=q.device)
    km = torch.empty((b, hk, d), dtype=k.dtype, device=k.device)
    n_stat = nk if _n_stat is None else max(nk, int(_n_stat))
    npad = (n_stat + 63) // 64 * 64
    if tensor_layout == "HND":
        v_fp8 = torch.empty((b, hk, d, npad), dtype=torch.float8_e4m3fn, device=v.device)
        sb, sh, sd = v_fp8.stride(0), v_fp8.stride(1), v_fp8.stride(2)
    else:
        v_fp8 = torch.empty((b, d, hk, npad), dtype=torch.float8_e4m3fn, device=v.device)
        sb, sh, sd = v_fp8.stride(0), v_fp8.stride(2), v_fp8.stride(1)
    v_scale = torch.empty((b, hk, d), dtype=torch.float32, device=v.device)
    with _on_device(q.device):
        _lib.check(lib.sage_prepass_qkv_fp8(q4, k4, qo4, ko4, q_scale.data_ptr(), k_scale.data_ptr(), km.data_ptr(),
                                            v4, v_fp8.data_ptr(), sb, sh, sd, v_scale.data_ptr(), b, hq, hk, nq, nk,
                                            n_stat, d, _dtype_code(q), per_thread, WARPQ, q_per_bh, k_per_bh,
                                            float(scale_max), d_og, _stream(q)))
    return q_int8, q_scale, k_int8, k_scale, km, v_fp8, v_scale


def quant_q_int8(q, qk_quant_gran="per_thread", WARPQ=32, BLKQ=128, tensor_layout="HND"):
    """The Q half of per_warp_int8 / per_thread_int8 alone."""
    _require_cuda(q)
    lib = _lib.load()
    b, hq, nq, d, q4 = _dims(q, tensor_layout)
    if qk_quant_gran == "per_warp":
        return _quant_block(q, None, WARPQ, None, tensor_layout,
                            scale_per_bh=((nq + BLKQ - 1) // BLKQ) * (BLKQ // WARPQ))
    q_int8 = torch.empty(q.shape, dtype=torch.int8, device=q.device)
    _, _, _, _, qo4 = _dims(q_int8, tensor_layout)
    q_scale = torch.empty((b, hq, (nq + BLKQ - 1) // BLKQ * (BLKQ // WARPQ) * 8), dtype=torch.float32, device=q.device)
    with _on_device(q.device):
        _lib.check(lib.sage_quant_per_thread_int8(q4, None, qo4, q_scale.data_ptr(), b, hq, nq, d, _dtype_code(q),
                                                  0, WARPQ, _stream(q)))
    return q_int8, q_scale


def sub_mean(v, tensor_layout="HND", vm=None):
    """Reference: /root/reference/sageattention/quant.py:183-223 -> (v_smoothed fp16, vm input dtype).
    `vm` (optional, [B,H,D] in v's dtype) overrides the internally computed mean."""
    _require_cuda(v)
    lib = _lib.load()
    b, h, n, d, v4 = _dims(v, tensor_layout)
    if vm is None:
        vm = seq_mean(v, tensor_layout)
    vm = vm.reshape(b, h, d).contiguous()
    out = torch.empty(v.shape, dtype=torch.float16, device=v.device)
    _, _, _, _, o4 = _dims(out, tensor_layout)
    with _on_device(v.device):
        _lib.check(lib.sage_sub_mean(v4, vm.data_ptr(), o4, b, h, n, d, _dtype_code(v), _stream(v)))
    return out, vm


def per_channel_fp8(v, tensor_layout="HND", scale_max=448.0, smooth_v=True, _n_stat=None):
    """Reference: /root/reference/sageattention/quant.py:225-322 with upstream NVIDIA semantics
    (float8_e4m3fn; the fork's e4m3fnuz at quant.py:312 is an AMD format).
    -> (v_fp8 [B,H,D,Npad] ("HND") or [B,D,H,Npad] ("NHD"), v_scale [B,H,D] f32, vm [B,H,D] f32 | None)

    `_n_stat` (internal): behave as if v had been zero-padded along the sequence to that many
    tokens first, as core.py:822-828 does, without materialising the padded copy."""
    _require_cuda(v)
    lib = _lib.load()
    b, h, n, d, v4 = _dims(v, tensor_layout)
    n_stat = n if _n_stat is None else max(n, int(_n_stat))
    npad = (n_stat + 63) // 64 * 64
    if tensor_layout == "HND":
        v_fp8 = torch.empty((b, h, d, npad), dtype=torch.float8_e4m3fn, device=v.device)
        sb, sh, sd = v_fp8.stride(0), v_fp8.stride(1), v_fp8.stride(2)
    else:
        v_fp8 = torch.empty((b, d, h, npad), dtype=torch.float8_e4m3fn, device=v.device)
        sb, sh, sd = v_fp8.stride(0), v_fp8.stride(2), v_fp8.stride(1)
    v_scale = torch.empty((b, h, d), dtype=torch.float32, device=v.device)
    vm = torch.empty((b, h, d), dtype=torch.float32, device=v.device) if smooth_v else None
    ws = torch.empty(lib.sage_per_channel_fp8_workspace_bytes(b, h, n, d), dtype=torch.uint8, device=v.device)
    with _on_device(v.device):
        _lib.check(lib.sage_per_channel_fp8(v4, v_fp8.data_ptr(), sb, sh, sd, v_scale.data_ptr(),
                                            vm.data_ptr() if smooth_v else None, b, h, n, n_stat, d,
                                            _dtype_code(v), float(scale_max), 1 if smooth_v else 0,
                                            ws.data_ptr(), _stream(v)))
    return v_fp8, v_scale, vm


def per_block_int8_varlen(q, k, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, BLKQ=128, BLKK=64, sm_scale=None,
                          km=None):
    """Per-block INT8 quantisation of a packed variable-length batch in ONE launch, bit for bit the reference's
    Triton `per_block_int8` (/root/reference/sageattention/triton/quant_per_block_varlen.py:59-107): q [Tq, Hq, D] is
    pre-scaled by sm_scale * 1.44269504, blocks of BLKQ / BLKK rows restart at every sequence, scale = amax / 127 with
    no floor, round half away from zero.  `km` ([Hkv, D] or broadcastable, k's dtype): subtracted from k first, like
    the `k - km` of core.py:462-463, inside the same launch.

    Returns (q_int8, q_scale, k_int8, k_scale, cu_seqlens_q_scale, cu_seqlens_k_scale) as the reference does, with
    one difference that keeps the call free of host synchronisation: the scale arrays have ceil(T / BLK) + batch rows
    (an upper bound); rows [0, cu_seqlens_*_scale[-1]) are the reference's, the rest is never written.  The work-list
    prefix the attention kernel needs is returned as a 7th element."""
    _require_cuda(q, k)
    if BLKQ != 128 or BLKK != 64:
        raise ValueError("the sm_100a kernels use BLKQ = 128 and BLKK = 64")
    if q.dim() != 3 or k.dim() != 3:
        raise ValueError("q and k must be [tokens, heads, head_dim]")
    lib = _lib.load()
    tq, hq, d = q.shape
    tk, hk, _ = k.shape
    if d not in (64, 128):
        raise ValueError("head_dim must be 64 or 128 (pad first)")
    if q.stride(-1) != 1 or k.stride(-1) != 1:
        raise ValueError("last dim of q / k must be contiguous")
    if sm_scale is None:
        sm_scale = d ** -0.5
    nseq = cu_seqlens_q.numel() - 1
    cu_q = cu_seqlens_q.to(torch.int32).contiguous()
    cu_k = cu_seqlens_k.to(torch.int32).contiguous()
    dev = q.device
    q_int8 = torch.empty((tq, hq, d), dtype=torch.int8, device=dev)
    k_int8 = torch.empty((tk, hk, d), dtype=torch.int8, device=dev)
    q_scale = torch.empty(((tq + 127) // 128 + nseq, hq), dtype=torch.float32, device=dev)
    k_scale = torch.empty(((tk + 63) // 64 + nseq, hk), dtype=torch.float32, device=dev)
    cu_qs = torch.empty(nseq + 1, dtype=torch.int32, device=dev)
    cu_ks = torch.empty(nseq + 1, dtype=torch.int32, device=dev)
    cu_pairs = torch.empty(nseq + 1, dtype=torch.int32, device=dev)
    km_ptr = None
    if km is not None:
        km = km.reshape(hk, d).contiguous()
        if km.dtype != k.dtype:
            raise RuntimeError("km must have the same dtype as k")
        km_ptr = km.data_ptr()
    with _on_device(dev):
        _lib.check(lib.sage_quant_per_block_varlen_int8(
            q.data_ptr(), k.data_ptr(), km_ptr, q_int8.data_ptr(), k_int8.data_ptr(), q_scale.data_ptr(),
            k_scale.data_ptr(), cu_q.data_ptr(), cu_k.data_ptr(), cu_qs.data_ptr(), cu_ks.data_ptr(), cu_pairs.data_ptr(),
            nseq, hq, hk, d, _dtype_code(q), int(max_seqlen_q), int(max_seqlen_k), q.stride(0), q.stride(1), k.stride(0),
            k.stride(1), float(sm_scale) * 1.44269504, _stream(q)))
    return q_int8, q_scale, k_int8, k_scale, cu_qs, cu_ks, (cu_q, cu_k, cu_pairs)
