"""Public attention API - a drop-in for /root/reference/sageattention/core.py on B200.

    sageattn(q, k, v, tensor_layout="HND", is_causal=False, sm_scale=None, smooth_k=True, return_lse=False)
    sageattn_qk_int8_pv_fp16_cuda(...)      reference: core.py:480-653
    sageattn_qk_int8_pv_fp8_cuda(...)       reference: core.py:656-905 (upstream semantics)

Host code is Python/PyTorch (allocation, streams); all arithmetic runs in libsage_b200.so
(hand-written sm_100a CUDA, include/sage_b200.h).  There is no Triton path, no multi-arch dispatch
and no CPU fallback: a CPU tensor or a missing library raises.
"""
import warnings
from typing import Any, Optional

import ctypes

import torch

from . import _lib
from .quant import (_dims, _dtype_code, _on_device, _require_cuda, _stream, per_block_int8, per_block_int8_varlen, per_channel_fp8,
                    per_thread_int8,
                    per_warp_int8, prepass_qk_int8, prepass_qkv_fp8, quant_q_int8, seq_mean, smooth_quant_k_int8,
                    sub_mean)

_LOG2E = 1.44269504

# bench.py sets this to a list to collect (start, end) CUDA events around every attention-kernel
# launch (the roofline's "dominant kernel" duration); None = no events recorded.
KERNEL_EVENTS = None

# Upper bound on the attention kernel's persistent grid (None = one CTA per SM).  parallel.ulysses_sageattn sets
# it while head groups are pipelined, so that the NCCL all-to-all of one group finds free SMs next to the
# attention of another (sage_attn_args.max_ctas).
GRID_LIMIT = None


def _pad_head_dim(q, k, v):
    """Zero-pad head_dim to 64 / 128 like core.py:590-601."""
    d = q.size(-1)
    if d < 64:
        pad = 64 - d
    elif 64 < d < 128:
        pad = 128 - d
    elif d > 128:
        raise ValueError(f"Unsupported head_dim: {d}")
    else:
        return q, k, v
    f = torch.nn.functional.pad
    return f(q, (0, pad)), f(k, (0, pad)), f(v, (0, pad))


def _check_inputs(q, k, v, qk_quant_gran):
    assert q.is_cuda, "Input tensors must be on cuda."
    assert q.dtype in [torch.float16, torch.bfloat16], \
        "Input tensors must be in dtype of torch.float16 or torch.bfloat16"
    assert qk_quant_gran in ["per_warp", "per_thread"], "qk_quant_gran must be either 'per_warp' or 'per_thread'."
    assert q.device == k.device == v.device, "All tensors must be on the same device."
    assert q.dtype == k.dtype == v.dtype, "All tensors must have the same dtype."


def _lse_correction(q, km, tensor_layout):
    """q . km, the term the smoothing removes from every logit of a row (core.py:613-617).
    With GQA the reference's matmul only broadcasts when Hq == Hkv; here km is expanded per q head."""
    hk = km.size(1)
    hq = q.size(1) if tensor_layout == "HND" else q.size(2)
    kmq = km if hq == hk else km.repeat_interleave(hq // hk, dim=1)
    qh = q if tensor_layout == "HND" else q.transpose(1, 2)
    return torch.matmul(qh, kmq.unsqueeze(-1)).squeeze(-1).to(torch.float32)


def _fold_padding(head_dim, smooth_k, smooth_v):
    """head_dim not 64 / 128: the sm_100a kernels can take the tensors as they are (multiples of 8, default
    smooth_k path without smooth_v): the quantisers write codes of the padded width, TMA zero-fills V and clips
    O.  Otherwise the reference's F.pad copies (core.py:590-601) are made."""
    return head_dim not in (64, 128) and head_dim <= 128 and head_dim % 8 == 0 and smooth_k and not smooth_v


def _quantize_qk(q, k, smooth_k, qk_quant_gran, tensor_layout, warp_q, return_lse, fold=False, v_to_f16=None):
    """The quantisation prologue (core.py:611-632): Q codes, and for K either the fused
    mean + quantise launch (smooth_k) or the plain quantiser.  Returns the four quantiser outputs
    and the LSE correction (None unless smooth_k and return_lse).
    `v_to_f16` (smooth_k only): a bf16 V whose f16 copy (core.py:633) the same launch writes; it is then
    returned as a sixth element."""
    if smooth_k:
        # one cluster-launched grid for the whole prologue: K slices (mean + quantise, K read once
        # from HBM) and Q tiles side by side (measured at C2: per_thread 105 us against 41 + 72 us as
        # two launches, per_warp 89 us against 45 + 62 us)
        out = prepass_qk_int8(q, k, qk_quant_gran, WARPQ=warp_q, tensor_layout=tensor_layout, _pad_head_dim=fold,
                              v=v_to_f16)
        q_int8, q_scale, k_int8, k_scale, km = out[:5]
        corr = _lse_correction(q, km[..., :q.size(-1)], tensor_layout) if return_lse else None
        return (q_int8, q_scale, k_int8, k_scale, corr) + tuple(out[5:])
    if qk_quant_gran == "per_warp":
        out = per_warp_int8(q, k, None, tensor_layout=tensor_layout, BLKQ=128, WARPQ=warp_q, BLKK=64)
    else:
        out = per_thread_int8(q, k, None, tensor_layout=tensor_layout, BLKQ=128, WARPQ=warp_q, BLKK=64, WARPK=64)
    return (*out, None)


def _prepare_mask(attn_mask, q, b, hq, nq, nk):
    """attn_mask (bool, or q's dtype; broadcastable to [B, H, Nq, Nk] - core.py:249-251, 306-318) in the form the
    kernels read: 4-D, batch / head dims of size 1 kept as stride-0 broadcasts, rows of ceil128(Nk) elements, 16-byte
    aligned.  A mask that already has that form is used in place; otherwise ONE padded copy is made (of the
    un-broadcast mask: a [1, 1, Nq, Nk] mask is copied once, not B*H times)."""
    assert attn_mask.dtype == torch.bool or attn_mask.dtype == q.dtype, \
        "attn_mask must be of dtype bool or the same dtype as q."
    assert attn_mask.device == q.device, "All tensors must be on the same device."
    m = attn_mask
    while m.dim() < 4:
        m = m.unsqueeze(0)
    target = (b, hq, nq, nk)
    try:
        torch.broadcast_shapes(tuple(m.shape), target)
        ok = all(ms in (1, ts) for ms, ts in zip(m.shape, target)) and m.dim() == 4
    except RuntimeError:
        ok = False
    if not ok:
        raise AssertionError(f"attn_mask shape {tuple(attn_mask.shape)} cannot be broadcast to {target}")
    if m.size(2) == 1 or m.size(3) == 1:   # broadcast along the sequence dims: materialise those (small dims stay)
        m = m.expand(m.size(0), m.size(1), nq, nk)
    npad = (nk + 127) // 128 * 128
    esz = 1 if m.dtype == torch.bool else 2
    aligned = (m.stride(3) == 1 and m.stride(2) >= npad and (m.stride(2) * esz) % 16 == 0 and m.data_ptr() % 16 == 0
               and all(m.size(i) == 1 or (m.stride(i) * esz) % 16 == 0 for i in (0, 1)))
    if not aligned:
        buf = torch.zeros((m.size(0), m.size(1), nq, npad), dtype=m.dtype, device=m.device)
        buf[..., :nk].copy_(m)
        m = buf
    kind = 1 if m.dtype == torch.bool else (2 if m.dtype == torch.float16 else 3)
    sb = m.stride(0) if m.size(0) > 1 else 0
    sh = m.stride(1) if m.size(1) > 1 else 0
    return m, kind, sb, sh, m.stride(2)


def _launch(fp8, q_int8, k_int8, v, o, q_scale, k_scale, v_scale, v_mean, tensor_layout, is_causal, gran,
            warp_q, sm_scale, return_lse, npad_v=0, debug_s=None, mask=None, varlen=None, o_peers=None):
    """One call into sage_attn_i8_f8 / sage_attn_i8_f16; returns the base-2 LSE or None."""
    lib = _lib.load()
    b, hq, nq, d, q4 = _dims(q_int8, tensor_layout)
    _, hk, nk, _, k4 = _dims(k_int8, tensor_layout)
    _, _, _, _, o4 = _dims(o, tensor_layout)
    if fp8:
        # v_fp8 is [B,H,D,Npad] (HND) or [B,D,H,Npad] (NHD); rows of Npad bytes
        if tensor_layout == "HND":
            v4 = _lib.Tensor4(v.data_ptr(), v.stride(0), v.stride(1), v.stride(2))
        else:
            v4 = _lib.Tensor4(v.data_ptr(), v.stride(0), v.stride(2), v.stride(1))
        npad_v = v.size(-1)
    else:
        _, _, _, _, v4 = _dims(v, tensor_layout)
    lse = torch.empty((b, hq, nq), dtype=torch.float32, device=o.device) if return_lse else None
    a = _lib.AttnArgs()
    a.q8, a.k8, a.v, a.o = q4, k4, v4, o4
    a.q_scale, a.k_scale = q_scale.data_ptr(), k_scale.data_ptr()
    a.v_scale = v_scale.data_ptr() if v_scale is not None else None
    a.v_mean = v_mean.data_ptr() if v_mean is not None else None
    a.lse = lse.data_ptr() if lse is not None else None
    a.B, a.Hq, a.Hkv, a.Nq, a.Nk, a.D = b, hq, hk, nq, nk, d
    a.npad_v = npad_v
    a.gran = gran
    a.warp_q = warp_q
    a.is_causal = 1 if is_causal else 0
    a.out_dtype = _dtype_code(o)
    a.sm_scale = float(sm_scale)
    a.debug_s = debug_s.data_ptr() if debug_s is not None else None
    a.d_og = o.size(-1) if o.size(-1) != d else 0
    a.max_ctas = int(GRID_LIMIT) if GRID_LIMIT else 0
    if mask is not None:
        mt, a.mask_kind, a.mask_stride_b, a.mask_stride_h, a.mask_stride_m = mask
        a.mask = mt.data_ptr()
    if varlen is not None:
        cu_q, cu_k, cu_qs, cu_ks, cu_pairs = varlen
        a.cu_seqlens_q, a.cu_seqlens_k = cu_q.data_ptr(), cu_k.data_ptr()
        a.cu_seqlens_q_scale, a.cu_seqlens_k_scale, a.cu_pairs = cu_qs.data_ptr(), cu_ks.data_ptr(), cu_pairs.data_ptr()
        a.nseq = cu_q.numel() - 1
    if o_peers is not None:
        # sequence-parallel output scatter: (addresses of the destinations' output rows, rows per destination,
        # head stride, row stride, device scratch for the tensor maps) - include/sage_b200.h, o_peer_*
        ptrs, rows, stride_h, stride_n, scratch = o_peers
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        a.o_peer_ptrs = ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p))
        a.o_peer_count, a.o_peer_rows = len(ptrs), rows
        a.o_peer_stride_h, a.o_peer_stride_n = stride_h, stride_n
        a.o_peer_maps = scratch.data_ptr()
    fn = lib.sage_attn_i8_f8 if fp8 else lib.sage_attn_i8_f16
    with _on_device(o.device):
        if KERNEL_EVENTS is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(fn(a, _stream(o)))
            e1.record()
            KERNEL_EVENTS.append((e0, e1))
        else:
            _lib.check(fn(a, _stream(o)))
    return lse


# One foreign call + one workspace per operator call (sage_sageattn_fwd) on the default paths; False = the two-call
# path with one tensor per intermediate (kept for the non-default options and as the reference the fast path is
# tested against: results are bit-identical up to 65 536 kv rows; above, the library issues the prologue as flat
# launches - faster for a few long slices - whose K mean can differ from the cluster kernel's in the last bit).
FUSED_CALL = True
_WS_BYTES = {}


def _fused_forward(q, k, v, tensor_layout, is_causal, qk_quant_gran, warp_q, sm_scale, pv_fp8, v_scale_max, return_lse):
    """sage_sageattn_fwd: prologue (smooth_k, no smooth_v) + attention from ONE ctypes crossing; the host side is two
    allocations (workspace, output; two more with return_lse).  Host time per call about 25 us against about 75 us for
    the two-call path (tools/host_overhead.py)."""
    lib = _lib.load()
    hnd = tensor_layout == "HND"
    if hnd:
        b, hq, nq, d_og = q.shape
        hk, nk = k.shape[1], k.shape[2]
        hi, ni = 1, 2
    elif tensor_layout == "NHD":
        b, nq, hq, d_og = q.shape
        nk, hk = k.shape[1], k.shape[2]
        hi, ni = 2, 1
    else:
        raise ValueError(f"Unknown tensor layout: {tensor_layout}")
    d = 64 if d_og <= 64 else 128
    dev = q.device
    T4 = _lib.Tensor4
    o = torch.empty(q.shape, dtype=q.dtype, device=dev)
    a = _lib.FwdArgs()
    a.q = T4(q.data_ptr(), q.stride(0), q.stride(hi), q.stride(ni))
    a.k = T4(k.data_ptr(), k.stride(0), k.stride(hi), k.stride(ni))
    a.v = T4(v.data_ptr(), v.stride(0), v.stride(hi), v.stride(ni))
    a.o = T4(o.data_ptr(), o.stride(0), o.stride(hi), o.stride(ni))
    a.B, a.Hq, a.Hkv, a.Nq, a.Nk, a.D = b, hq, hk, nq, nk, d
    a.d_og = d_og if d_og != d else 0
    a.dtype = _dtype_code(q)
    a.layout_hnd = 1 if hnd else 0
    a.pv_fp8 = 1 if pv_fp8 else 0
    a.per_thread = 1 if qk_quant_gran == "per_thread" else 0
    a.warp_q = warp_q
    a.is_causal = 1 if is_causal else 0
    a.max_ctas = int(GRID_LIMIT) if GRID_LIMIT else 0
    a.sm_scale = float(sm_scale)
    a.v_scale_max = float(v_scale_max)
    key = (b, hq, hk, nq, nk, d, d_og, a.dtype, a.pv_fp8, a.per_thread, warp_q)
    nbytes = _WS_BYTES.get(key)
    if nbytes is None:
        nbytes = lib.sage_sageattn_fwd_workspace_bytes(a)
        if nbytes == 0:
            _lib.check(-1)
        _WS_BYTES[key] = nbytes
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    a.workspace, a.workspace_bytes = ws.data_ptr(), nbytes
    lse = km = None
    if return_lse:
        lse = torch.empty((b, hq, nq), dtype=torch.float32, device=dev)
        km = torch.empty((b, hk, d), dtype=k.dtype, device=dev)
        a.lse, a.km_out = lse.data_ptr(), km.data_ptr()
    ev = None
    if KERNEL_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        for e in ev:
            e.record()   # creates the underlying cudaEvent_t; the library records it again around the attention launch
        a.ev_start, a.ev_stop = ev[0].cuda_event, ev[1].cuda_event
    with _on_device(dev):
        _lib.check(lib.sage_sageattn_fwd(a, _stream(q)))
    if ev is not None:
        KERNEL_EVENTS.append(ev)
    corr = _lse_correction(q, km[..., :d_og], tensor_layout) if return_lse else None
    return _finish(o, lse, corr, sm_scale, d_og, return_lse)


def _finish(o, lse, lse_correction, sm_scale, head_dim_og, return_lse):
    o = o[..., :head_dim_og]
    if not return_lse:
        return o
    lse = lse / _LOG2E
    if lse_correction is not None:
        lse = lse + lse_correction * sm_scale
    return o, lse


@torch.compiler.disable
def sageattn_qk_int8_pv_fp16_cuda(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    tensor_layout: str = "HND",
    is_causal: bool = False,
    qk_quant_gran: str = "per_thread",
    sm_scale: Optional[float] = None,
    pv_accum_dtype: str = "fp32",
    smooth_k: bool = True,
    smooth_v: bool = False,
    return_lse: bool = False,
    **kwargs: Any,
) -> torch.Tensor:
    """INT8 QK^T + FP16 P.V (reference core.py:480-653).

    pv_accum_dtype in {"fp32", "fp16", "fp16+fp32"} is accepted for compatibility; on sm_100a the
    P.V accumulator always lives in TMEM as fp32 (the reference's fp16 accumulators exist only
    because mma.sync fp16 accumulation was faster on sm80), so all three give fp32-accumulated
    results.  `smooth_v` is honoured for "fp16" as in the reference and ignored (with the same
    warning) otherwise.  Extra kwargs (attn_mask, dropout_p, ...) are ignored like the reference.
    """
    _check_inputs(q, k, v, qk_quant_gran)
    if pv_accum_dtype not in ("fp32", "fp16", "fp16+fp32"):
        raise ValueError(f"Unsupported pv_accum_dtype: {pv_accum_dtype}")
    dtype = q.dtype
    head_dim_og = q.size(-1)
    if pv_accum_dtype in ["fp32", "fp16+fp32"] and smooth_v:
        warnings.warn(f"pv_accum_dtype is {pv_accum_dtype}, smooth_v will be ignored.")
        smooth_v = False
    fold = _fold_padding(head_dim_og, smooth_k, smooth_v)
    if not fold:
        q, k, v = _pad_head_dim(q, k, v)
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "Last dim of qkv must be contiguous."
    if sm_scale is None:
        sm_scale = head_dim_og ** -0.5
    warp_q = 16 if (head_dim_og > 64 and pv_accum_dtype == "fp16+fp32") else 32
    if FUSED_CALL and smooth_k and not smooth_v and q.dim() == 4 and v.shape == k.shape and \
            (fold or head_dim_og in (64, 128)):
        return _fused_forward(q, k, v, tensor_layout, is_causal, qk_quant_gran, warp_q, sm_scale, False, 0.0, return_lse)
    # a bf16 V needs an f16 copy (core.py:633): on the default path it comes out of the prologue launch
    fuse_v = smooth_k and not smooth_v and v.dtype == torch.bfloat16 and v.shape == k.shape
    qk = _quantize_qk(q, k, smooth_k, qk_quant_gran, tensor_layout, warp_q, return_lse, fold,
                      v_to_f16=v if fuse_v else None)
    q_int8, q_scale, k_int8, k_scale, lse_correction = qk[:5]
    o = torch.empty(q.size(), dtype=dtype, device=q.device)
    v_mean = None
    if smooth_v:
        v16, vm = sub_mean(v, tensor_layout=tensor_layout)
        v_mean = vm.to(torch.float32)
    elif fuse_v:
        v16 = qk[5]
    else:
        v16 = v.to(torch.float16)
    gran = _lib.GRAN_PER_THREAD if qk_quant_gran == "per_thread" else _lib.GRAN_PER_WARP
    lse = _launch(False, q_int8, k_int8, v16, o, q_scale, k_scale, None, v_mean, tensor_layout, is_causal, gran,
                  warp_q, sm_scale, return_lse)
    return _finish(o, lse, lse_correction, sm_scale, head_dim_og, return_lse)


@torch.compiler.disable
def sageattn_qk_int8_pv_fp8_cuda(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    tensor_layout: str = "HND",
    is_causal: bool = False,
    qk_quant_gran: str = "per_thread",
    sm_scale: Optional[float] = None,
    pv_accum_dtype: str = "fp32+fp16",
    smooth_k: bool = True,
    smooth_v: bool = False,
    return_lse: bool = False,
    **kwargs: Any,
) -> torch.Tensor:
    """INT8 QK^T + FP8(e4m3) P.V (reference core.py:656-905, upstream semantics).

    pv_accum_dtype in {"fp32", "fp32+fp32", "fp32+fp16"}: accumulation is always fp32 in TMEM;
    the string still selects the V quantisation range exactly as the reference does
    (scale_max 448, or 2.25 for "fp32+fp16", core.py:818-820) and whether smooth_v applies.
    """
    _check_inputs(q, k, v, qk_quant_gran)
    if pv_accum_dtype not in ("fp32", "fp32+fp32", "fp32+fp16"):
        raise ValueError(f"Unsupported pv_accum_dtype: {pv_accum_dtype}")
    dtype = q.dtype
    head_dim_og = q.size(-1)
    if pv_accum_dtype == "fp32+fp32" and smooth_v:
        warnings.warn("pv_accum_dtype is 'fp32+fp32', smooth_v will be ignored.")
        smooth_v = False
    if pv_accum_dtype == "fp32+fp16" and smooth_v:
        warnings.warn("pv_accum_dtype is 'fp32+fp16', smooth_v will be ignored.")
        smooth_v = False
    fold = _fold_padding(head_dim_og, smooth_k, smooth_v) and k.shape == v.shape
    if not fold:
        q, k, v = _pad_head_dim(q, k, v)
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "Last dim of qkv must be contiguous."
    if sm_scale is None:
        sm_scale = head_dim_og ** -0.5
    quant_v_scale_max = 2.25 if pv_accum_dtype == "fp32+fp16" else 448.0
    if FUSED_CALL and smooth_k and not smooth_v and q.dim() == 4 and k.shape == v.shape and \
            (fold or head_dim_og in (64, 128)):
        return _fused_forward(q, k, v, tensor_layout, is_causal, qk_quant_gran, 32, sm_scale, True, quant_v_scale_max,
                              return_lse)
    o = torch.empty(q.size(), dtype=dtype, device=q.device)
    kv_len = k.size(2) if tensor_layout == "HND" else k.size(1)
    # the reference zero-pads V to a multiple of 128 tokens first (core.py:822-828); the kernels
    # reproduce that padding's effect on Npad / the smooth_v mean without the extra copy
    n_stat = (kv_len + 127) // 128 * 128
    if smooth_k and not smooth_v and k.shape == v.shape:
        # the whole prologue in ONE launch: K slices (mean + quantise), V slices (range + e4m3 transpose) and
        # Q tiles side by side (bit-identical to the separate launches; measured at C2 in DESIGN 3.2)
        q_int8, q_scale, k_int8, k_scale, km, v_fp8, v_scale = prepass_qkv_fp8(
            q, k, v, qk_quant_gran, WARPQ=32, tensor_layout=tensor_layout, scale_max=quant_v_scale_max, _n_stat=n_stat,
            _pad_head_dim=fold)
        lse_correction = _lse_correction(q, km[..., :q.size(-1)], tensor_layout) if return_lse else None
        vm = None
    else:
        q_int8, q_scale, k_int8, k_scale, lse_correction = _quantize_qk(q, k, smooth_k, qk_quant_gran, tensor_layout,
                                                                        32, return_lse)
        v_fp8, v_scale, vm = per_channel_fp8(v, tensor_layout=tensor_layout, scale_max=quant_v_scale_max,
                                             smooth_v=smooth_v, _n_stat=n_stat)
    gran = _lib.GRAN_PER_THREAD if qk_quant_gran == "per_thread" else _lib.GRAN_PER_WARP
    lse = _launch(True, q_int8, k_int8, v_fp8, o, q_scale, k_scale, v_scale, vm, tensor_layout, is_causal, gran,
                  32, sm_scale, return_lse)
    return _finish(o, lse, lse_correction, sm_scale, head_dim_og, return_lse)


def sageattn(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    tensor_layout: str = "HND",
    is_causal: bool = False,
    sm_scale: Optional[float] = None,
    smooth_k: bool = True,
    return_lse: bool = False,
    **kwargs: Any,
):
    """Reference: core.py:80-158.  The reference picks a kernel per GPU architecture (and the fork
    hard-wires the fp8 path, core.py:144); this package targets sm_100a only, so `sageattn` is the
    INT8-QK / FP8-PV tcgen05 kernel with fp32 accumulation.  `smooth_k` is accepted and forwarded
    (the reference drops it inside **kwargs)."""
    return sageattn_qk_int8_pv_fp8_cuda(q, k, v, tensor_layout=tensor_layout, is_causal=is_causal,
                                        sm_scale=sm_scale, pv_accum_dtype="fp32+fp32", smooth_k=smooth_k,
                                        return_lse=return_lse)


@torch.compiler.disable
def sageattn_qk_int8_pv_fp16_triton(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    tensor_layout: str = "HND",
    quantization_backend: str = "triton",
    is_causal: bool = False,
    attn_mask: Optional[torch.Tensor] = None,
    sm_scale: Optional[float] = None,
    smooth_k: bool = True,
    return_lse: bool = False,
    **kwargs: Any,
) -> torch.Tensor:
    """The reference's Triton entry point (core.py:160-361; exported by its __init__ because diffusers imports it),
    and the one that takes an `attn_mask`.  There is no Triton here: the call runs on the same sm_100a kernels as
    sageattn_qk_int8_pv_fp16_cuda (INT8 QK^T with per-thread scales, f16 P.V, fp32 accumulation in TMEM).

    attn_mask: bool (True = attend) or q's dtype (ADDED to the base-2 logits as it is - the reference does not scale
    it by log2 e, triton/attn_qk_int8_per_thread.py:68-72), broadcastable to [B, H, Nq, Nk]; not with is_causal.
    Masked keys contribute exactly 0; a query row with every key masked gives zeros (the reference's finite -1e6
    leaves an unmasked softmax over the partially masked blocks there, SDPA gives NaN).
    quantization_backend ("triton" | "cuda") is accepted for compatibility; both use per-thread quantisation here
    (the reference feeds per-thread or per-block scales to the same per-thread kernel, core.py:295-299)."""
    dtype = q.dtype
    assert q.is_cuda, "Input tensors must be on cuda."
    assert dtype in [torch.float16, torch.bfloat16], \
        "Input tensors must be in dtype of torch.float16 or torch.bfloat16"
    assert q.device == k.device == v.device, "All tensors must be on the same device."
    assert q.dtype == k.dtype == v.dtype, "All tensors must have the same dtype."
    if quantization_backend not in ("triton", "cuda"):
        raise ValueError(f"Unsupported quantization backend: {quantization_backend}")
    if attn_mask is None:
        return sageattn_qk_int8_pv_fp16_cuda(q, k, v, tensor_layout=tensor_layout, is_causal=is_causal,
                                             qk_quant_gran="per_thread", sm_scale=sm_scale, pv_accum_dtype="fp32",
                                             smooth_k=smooth_k, return_lse=return_lse)
    assert not is_causal, "Mask should be None for causal attention."
    head_dim_og = q.size(-1)
    q, k, v = _pad_head_dim(q, k, v)
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "Last dim of qkv must be contiguous."
    if sm_scale is None:
        sm_scale = head_dim_og ** -0.5
    b, hq, nq, _, _ = _dims(q, tensor_layout)
    _, _, nk, _, _ = _dims(k, tensor_layout)
    mask = _prepare_mask(attn_mask, q, b, hq, nq, nk)
    q_int8, q_scale, k_int8, k_scale, lse_correction = _quantize_qk(q, k, smooth_k, "per_thread", tensor_layout, 32,
                                                                    return_lse)
    o = torch.empty(q.size(), dtype=dtype, device=q.device)
    lse = _launch(False, q_int8, k_int8, v.to(torch.float16), o, q_scale, k_scale, None, None, tensor_layout, False,
                  _lib.GRAN_PER_THREAD, 32, sm_scale, return_lse, mask=mask)
    return _finish(o, lse, lse_correction, sm_scale, head_dim_og, return_lse)


@torch.compiler.disable
def sageattn_varlen(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    cu_seqlens_q: torch.Tensor,
    cu_seqlens_k: torch.Tensor,
    max_seqlen_q: int,
    max_seqlen_k: int,
    is_causal: bool = False,
    sm_scale: Optional[float] = None,
    smooth_k: bool = True,
    **kwargs: Any,
) -> torch.Tensor:
    """Packed variable-length batches (reference core.py:363-477): q [Tq, Hq, D], k / v [Tk, Hkv, D],
    cu_seqlens_* [batch + 1]; returns [Tq, Hq, D].

    Same scheme as the reference: K is smoothed by its mean over ALL tokens of the pack (core.py:462), Q / K are
    quantised per block (128 / 64 rows, Q pre-scaled by sm_scale * log2 e, blocks restarting at every sequence) with
    the arithmetic of the reference's Triton quantiser bit for bit (quant.per_block_int8_varlen), V is f16, P.V
    accumulates in fp32.  Three launches for the whole pack whatever its make-up - the pack-wide K mean, ONE quantiser
    launch for Q and K, ONE attention launch whose persistent work list is derived from cu_seqlens on the device - and
    no host synchronisation: the call can be captured in a CUDA graph."""
    dtype = q.dtype
    assert q.is_cuda, "Input tensors must be on cuda."
    assert dtype in [torch.float16, torch.bfloat16], \
        "Input tensors must be in dtype of torch.float16 or torch.bfloat16"
    assert q.device == k.device == v.device, "All tensors must be on the same device."
    assert q.dtype == k.dtype == v.dtype, "All tensors must have the same dtype."
    assert q.dim() == 3 and k.dim() == 3 and v.dim() == 3, "q, k, v must be [tokens, heads, head_dim]"
    head_dim_og = q.size(-1)
    q, k, v = _pad_head_dim(q, k, v)
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "Last dim of qkv must be contiguous."
    assert cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous(), \
        "cu_seqlens_q and cu_seqlens_k must be contiguous."
    assert cu_seqlens_q.numel() == cu_seqlens_k.numel() and cu_seqlens_q.numel() >= 2
    if sm_scale is None:
        sm_scale = head_dim_og ** -0.5
    tq, hq, d = q.shape
    tk, hk, _ = k.shape
    o = torch.empty(q.shape, dtype=dtype, device=q.device)
    if tq == 0:
        return o[..., :head_dim_og]
    if tk == 0:
        return o.zero_()[..., :head_dim_og]
    v16 = v.to(torch.float16)
    # mean over all sequences of the pack, in the input dtype (core.py:462); the subtraction happens in the quantiser
    km = seq_mean(k.unsqueeze(0), tensor_layout="NHD") if smooth_k else None
    q_int8, q_scale, k_int8, k_scale, cu_qs, cu_ks, (cu_q, cu_k, cu_pairs) = per_block_int8_varlen(
        q, k, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, sm_scale=sm_scale, km=km)
    # Q already carries sm_scale * log2 e: the kernel's own factor must be 1
    _launch(False, q_int8.unsqueeze(0), k_int8.unsqueeze(0), v16.unsqueeze(0), o.unsqueeze(0), q_scale, k_scale, None, None,
            "NHD", is_causal, _lib.GRAN_PER_BLOCK, 32, 1.0 / 1.44269504088896340736, False,
            varlen=(cu_q, cu_k, cu_qs, cu_ks, cu_pairs))
    return o[..., :head_dim_og]
