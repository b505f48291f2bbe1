"""ctypes binding of libsage_b200.so (include/sage_b200.h).

There is no CPU fallback: if the library is missing it is built (nvcc), and if that fails the
import raises.  Every call checks the int status and raises ValueError (bad argument) or
RuntimeError (CUDA error) with the library's message, mirroring the reference's pybind
exception mapping (std::invalid_argument -> ValueError, TORCH_CHECK -> RuntimeError;
/root/reference/csrc/dispatch_utils.h:31-35, /root/reference/csrc/utils.cuh:20-37).
"""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("SAGE_B200_LIB") or os.path.join(_PKG, "libsage_b200.so")  # env: kernel-variant A/B runs

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_f32 = ctypes.c_float
c_vp = ctypes.c_void_p


class Tensor4(ctypes.Structure):
    _fields_ = [("ptr", c_vp), ("stride_b", c_i64), ("stride_h", c_i64), ("stride_n", c_i64)]


class AttnArgs(ctypes.Structure):
    _fields_ = [
        ("q8", Tensor4), ("k8", Tensor4), ("v", Tensor4), ("o", Tensor4),
        ("q_scale", c_vp), ("k_scale", c_vp), ("v_scale", c_vp), ("v_mean", c_vp), ("lse", c_vp),
        ("B", c_int), ("Hq", c_int), ("Hkv", c_int), ("Nq", c_int), ("Nk", c_int), ("D", c_int),
        ("npad_v", c_int), ("gran", c_int), ("warp_q", c_int), ("is_causal", c_int),
        ("out_dtype", c_int), ("sm_scale", c_f32), ("debug_s", c_vp), ("d_og", c_int), ("max_ctas", c_int),
        ("mask", c_vp), ("mask_kind", c_int), ("mask_stride_b", c_i64), ("mask_stride_h", c_i64), ("mask_stride_m", c_i64),
        ("cu_seqlens_q", c_vp), ("cu_seqlens_k", c_vp), ("cu_seqlens_q_scale", c_vp), ("cu_seqlens_k_scale", c_vp),
        ("cu_pairs", c_vp), ("nseq", c_int),
        ("o_peer_ptrs", ctypes.POINTER(c_vp)), ("o_peer_count", c_int), ("o_peer_rows", c_int),
        ("o_peer_stride_h", c_i64), ("o_peer_stride_n", c_i64), ("o_peer_maps", c_vp),
    ]


class FwdArgs(ctypes.Structure):
    _fields_ = [
        ("q", Tensor4), ("k", Tensor4), ("v", Tensor4), ("o", Tensor4),
        ("workspace", c_vp), ("workspace_bytes", ctypes.c_size_t), ("km_out", c_vp), ("lse", c_vp),
        ("ev_start", c_vp), ("ev_stop", c_vp),
        ("B", c_int), ("Hq", c_int), ("Hkv", c_int), ("Nq", c_int), ("Nk", c_int), ("D", c_int), ("d_og", c_int),
        ("dtype", c_int), ("layout_hnd", c_int), ("pv_fp8", c_int), ("per_thread", c_int), ("warp_q", c_int),
        ("is_causal", c_int), ("max_ctas", c_int), ("sm_scale", c_f32), ("v_scale_max", c_f32),
    ]


# name -> (restype, argtypes); the single source of truth for the exported surface
SIGNATURES = {
    "sage_last_error": (ctypes.c_char_p, []),
    "sage_version": (c_int, []),
    "sage_source_hash": (ctypes.c_char_p, []),
    "sage_seq_mean_workspace_bytes": (ctypes.c_size_t, [c_int] * 4),
    "sage_seq_mean": (c_int, [Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "sage_seq_mean_static": (c_int, [Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp]),
    "sage_quant_per_block_varlen_int8": (c_int, [c_vp] * 12 + [c_int] * 7 + [c_i64] * 4 + [c_f32, c_vp]),
    "sage_quant_int8": (c_int, [Tensor4, c_vp, Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int, c_int,
                                c_int, c_int, c_f32, c_vp]),
    "sage_quant_per_thread_int8": (c_int, [Tensor4, c_vp, Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int,
                                           c_int, c_int, c_vp]),
    "sage_quant_per_thread_int8_strided": (c_int, [Tensor4, c_vp, Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int,
                                                   c_int, c_int, c_int, c_vp]),
    "sage_smooth_quant_k_int8": (c_int, [Tensor4, c_vp, Tensor4, c_vp, c_int, c_int, c_int, c_int, c_int,
                                         c_int, c_int, c_vp]),
    "sage_prepass_qk_int8": (c_int, [Tensor4, Tensor4, Tensor4, Tensor4, c_vp, c_vp, c_vp] + [c_int] * 12 + [c_vp]),
    "sage_prepass_qk_int8_v16": (c_int, [Tensor4, Tensor4, Tensor4, Tensor4, c_vp, c_vp, c_vp, Tensor4, Tensor4]
                                 + [c_int] * 12 + [c_vp]),
    "sage_prepass_qkv_fp8": (c_int, [Tensor4, Tensor4, Tensor4, Tensor4, c_vp, c_vp, c_vp, Tensor4, c_vp, c_i64, c_i64,
                                     c_i64, c_vp] + [c_int] * 12 + [c_f32, c_int, c_vp]),
    "sage_sub_mean": (c_int, [Tensor4, c_vp, Tensor4, c_int, c_int, c_int, c_int, c_int, c_vp]),
    "sage_per_channel_fp8_workspace_bytes": (ctypes.c_size_t, [c_int] * 4),
    "sage_per_channel_fp8": (c_int, [Tensor4, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_int, c_int, c_int,
                                     c_int, c_int, c_int, c_f32, c_int, c_vp, c_vp]),
    "sage_per_channel_fp8_flat": (c_int, [Tensor4, c_vp, c_i64, c_i64, c_i64, c_vp, c_int, c_int, c_int, c_int, c_int, c_int,
                                          c_f32, c_vp, c_vp]),
    "sage_col_sum_absmax": (c_int, [Tensor4, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "sage_v_fp8_from_absmax": (c_int, [Tensor4, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int,
                                       c_int, c_f32, c_vp]),
    "sage_quant_per_thread_int8_peers": (c_int, [Tensor4, c_vp, ctypes.POINTER(c_vp), c_i64, c_i64, ctypes.POINTER(c_vp),
                                                 c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp]),
    "sage_v_fp8_from_absmax_peers": (c_int, [Tensor4, ctypes.POINTER(c_vp), c_i64, c_i64, c_vp, c_vp, c_int, c_int, c_int,
                                             c_int, c_int, c_int, c_f32, c_vp]),
    "sage_attn_i8_f16": (c_int, [ctypes.POINTER(AttnArgs), c_vp]),
    "sage_attn_i8_f8": (c_int, [ctypes.POINTER(AttnArgs), c_vp]),
    "sage_sageattn_fwd_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(FwdArgs)]),
    "sage_sageattn_fwd": (c_int, [ctypes.POINTER(FwdArgs), c_vp]),
}

_lib = None

# kernels launched by one call of each entry point (cudaMemsetAsync is not counted); LAUNCHES is
# bumped on every call so bench.py can report how many of OUR kernels ran in its timed region.
KERNELS_PER_CALL = {
    "sage_seq_mean": 1, "sage_seq_mean_static": 1, "sage_quant_per_block_varlen_int8": 1, "sage_quant_int8": 1, "sage_quant_per_thread_int8": 1, "sage_sub_mean": 1,
    "sage_smooth_quant_k_int8": 1, "sage_prepass_qk_int8": 1, "sage_prepass_qk_int8_v16": 1, "sage_prepass_qkv_fp8": 1,
    # smooth_v: statistics + conversion; otherwise one cluster launch
    "sage_per_channel_fp8": (lambda a: 2 if a[14] else 1),
    "sage_attn_i8_f16": 1, "sage_attn_i8_f8": 1,
    "sage_per_channel_fp8_flat": 2, "sage_col_sum_absmax": 1, "sage_quant_per_thread_int8_strided": 1,
    "sage_quant_per_thread_int8_peers": 1, "sage_v_fp8_from_absmax_peers": 1, "sage_v_fp8_from_absmax": 1,
    # prologue + attention; above 65 536 kv rows the prologue is flat launches (K mean, Q, K, V range, V conversion /
    # f16 copy of a bf16 V)
    "sage_sageattn_fwd": (lambda a: 2 if a[0].Nk <= 65536 else (6 if a[0].pv_fp8 else (5 if a[0].dtype == DTYPE_BF16 else 4))),
}
LAUNCHES = 0


class _Counted:
    """ctypes function wrapper that counts kernel launches."""

    def __init__(self, fn, n):
        self._fn, self._n = fn, n

    def __call__(self, *a):
        global LAUNCHES
        LAUNCHES += self._n(a) if callable(self._n) else self._n
        return self._fn(*a)


class _LibView:
    def __init__(self, lib):
        self._lib = lib
        for name in SIGNATURES:
            fn = getattr(lib, name)
            n = KERNELS_PER_CALL.get(name)
            setattr(self, name, _Counted(fn, n) if n else fn)


def load():
    """Load (building first if needed) and return the ctypes library handle."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        from . import build as _build
        _build.build()
    try:
        lib = _open(_LIB_PATH)
    except AttributeError:
        # a binary from before a symbol was added: same remedy as a hash mismatch
        if os.environ.get("SAGE_B200_LIB"):
            raise
        from . import build as _build
        _build.build(force=True)
        lib = _open(_LIB_PATH)
    if not os.environ.get("SAGE_B200_LIB"):
        # a binary built from other sources than the ones next to it (edited .cu / .h, an old .so that travelled): rebuild
        # when a compiler is at hand, else say so - never use it silently (round-1 ADVICE)
        try:
            from . import build as _build
            want = _build.source_hash()
            have = lib.sage_source_hash().decode() if hasattr(lib, "sage_source_hash") else "missing"
            if have != want:
                import shutil
                import warnings
                if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
                    _build.build()
                    lib = _open(_LIB_PATH)
                else:
                    warnings.warn(f"libsage_b200.so was built from other sources (hash {have[:12]}, sources {want[:12]}) and "
                                  "no nvcc is available to rebuild it")
        except OSError:
            pass  # sources not shipped with the package: nothing to compare with
    _lib = _LibView(lib)
    return _lib


def _open(path):
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


def check(rc):
    if rc == 0:
        return
    msg = load().sage_last_error().decode("utf-8", "replace")
    if rc < 0:
        raise ValueError(msg)
    raise RuntimeError(f"CUDA error {rc}: {msg}")


DTYPE_F16, DTYPE_BF16 = 0, 1
GRAN_PER_BLOCK, GRAN_PER_WARP, GRAN_PER_THREAD = 1, 2, 3
