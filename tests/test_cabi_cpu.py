"""CPU (`-m "not gpu"`): the C-ABI library builds, loads and exports every symbol that
include/sage_b200.h declares; argument validation (which runs before any CUDA call) returns the
documented status codes; the Python mirror keeps the reference's signatures and error types."""
import ctypes
import inspect
import os
import re

import pytest
import torch

import sageattention_b200 as sb
from sageattention_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "sage_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sage_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = ctypes.CDLL(os.path.join(ROOT, "sageattention_b200", "libsage_b200.so"))
    syms = _declared_symbols()
    assert len(syms) >= 11
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/sage_b200.h but not exported"
    assert set(_lib.SIGNATURES) == set(syms), "ctypes table and header disagree"


def test_version_and_error_string():
    lib = _lib.load()
    assert lib.sage_version() >= 100
    assert isinstance(lib.sage_last_error(), bytes)


def test_argument_validation_without_gpu():
    lib = _lib.load()
    t = _lib.Tensor4(16, 0, 0, 128)
    # bad head_dim -> negative status + message, before any CUDA call
    rc = lib.sage_quant_int8(t, None, t, 16, 1, 1, 64, 100, 0, 32, 2, 0, 0.0, None)
    assert rc < 0 and b"head_dim" in lib.sage_last_error()
    rc = lib.sage_quant_int8(t, None, t, 16, 1, 1, 64, 128, 7, 32, 2, 0, 0.0, None)
    assert rc < 0 and b"dtype" in lib.sage_last_error()
    rc = lib.sage_quant_int8(t, None, t, 16, 1, 1, 64, 128, 0, 48, 2, 0, 0.0, None)
    assert rc < 0 and b"block" in lib.sage_last_error().lower()
    a = _lib.AttnArgs()
    a.D = 96
    assert lib.sage_attn_i8_f8(ctypes.byref(a), None) < 0
    # the real head_dim of the one-launch prologues must be a multiple of 8 and <= the padded one
    rc = lib.sage_prepass_qk_int8(t, t, t, t, 16, 16, 16, 1, 1, 1, 64, 64, 128, 0, 1, 32, 16, 4, 100, None)
    assert rc < 0 and b"d_og" in lib.sage_last_error()
    rc = lib.sage_prepass_qkv_fp8(t, t, t, t, 16, 16, 16, t, 16, 0, 0, 0, 16, 1, 1, 1, 64, 64, 64, 64, 0, 1, 32, 16, 4,
                                  448.0, 72, None)
    assert rc < 0 and b"d_og" in lib.sage_last_error()
    # the sequence-parallel entry points: destination tables of at most 16 ranks, scale rows wide enough for the call's
    # own entries, a scratch buffer for the output maps - all rejected before any CUDA call
    ptrs = (ctypes.c_void_p * 17)(*([4096] * 17))
    rc = lib.sage_quant_per_thread_int8_peers(t, None, ptrs, 128, 256, ptrs, 17, 1, 64, 128, 0, 0, 32, 64, None)
    assert rc < 0 and b"16" in lib.sage_last_error()
    rc = lib.sage_quant_per_thread_int8_peers(t, None, ptrs, 128, 256, ptrs, 2, 1, 256, 128, 0, 1, 64, 8, None)
    assert rc < 0 and b"scale_per_bh" in lib.sage_last_error()
    rc = lib.sage_quant_per_thread_int8_strided(t, None, t, 16, 1, 1, 256, 128, 0, 0, 32, 32, None)
    assert rc < 0 and b"scale_per_bh" in lib.sage_last_error()
    rc = lib.sage_v_fp8_from_absmax_peers(t, ptrs, 128, 100, 16, None, 2, 1, 64, 64, 128, 0, 448.0, None)
    assert rc < 0 and b"16-byte" in lib.sage_last_error()
    rc = lib.sage_col_sum_absmax(t, None, None, 1, 1, 64, 128, 0, 16, None)
    assert rc < 0 and b"null" in lib.sage_last_error()
    with pytest.raises(ValueError):
        _lib.check(-1)


def test_python_api_matches_reference_signatures():
    """Names, positional order and defaults of /root/reference/sageattention/core.py:80-89,480-493,
    656-669 and quant.py:23,106,183,225."""
    def params(f):
        return [(n, p.default) for n, p in inspect.signature(f).parameters.items() if p.kind != p.VAR_KEYWORD]
    E = inspect.Parameter.empty
    assert params(sb.sageattn)[:6] == [("q", E), ("k", E), ("v", E), ("tensor_layout", "HND"),
                                       ("is_causal", False), ("sm_scale", None)]
    assert dict(params(sb.sageattn))["return_lse"] is False
    assert params(sb.sageattn_qk_int8_pv_fp16_cuda) == [
        ("q", E), ("k", E), ("v", E), ("tensor_layout", "HND"), ("is_causal", False),
        ("qk_quant_gran", "per_thread"), ("sm_scale", None), ("pv_accum_dtype", "fp32"),
        ("smooth_k", True), ("smooth_v", False), ("return_lse", False)]
    assert params(sb.sageattn_qk_int8_pv_fp8_cuda) == [
        ("q", E), ("k", E), ("v", E), ("tensor_layout", "HND"), ("is_causal", False),
        ("qk_quant_gran", "per_thread"), ("sm_scale", None), ("pv_accum_dtype", "fp32+fp16"),
        ("smooth_k", True), ("smooth_v", False), ("return_lse", False)]
    # core.py:363-375
    assert params(sb.sageattn_varlen) == [
        ("q", E), ("k", E), ("v", E), ("cu_seqlens_q", E), ("cu_seqlens_k", E), ("max_seqlen_q", E),
        ("max_seqlen_k", E), ("is_causal", False), ("sm_scale", None), ("smooth_k", True)]
    assert params(sb.per_block_int8) == [("q", E), ("k", E), ("km", None), ("BLKQ", 128), ("BLKK", 64),
                                         ("sm_scale", None), ("tensor_layout", "HND")]
    assert params(sb.per_warp_int8) == [("q", E), ("k", E), ("km", None), ("BLKQ", 128), ("WARPQ", 32),
                                        ("BLKK", 64), ("tensor_layout", "HND")]
    assert params(sb.per_channel_fp8)[:4] == [("v", E), ("tensor_layout", "HND"), ("scale_max", 448.0),
                                              ("smooth_v", True)]


def test_no_cpu_fallback():
    q = torch.randn(1, 2, 64, 64, dtype=torch.float16)
    with pytest.raises(AssertionError):
        sb.sageattn(q, q, q)
    with pytest.raises(AssertionError):
        cu = torch.tensor([0, 64], dtype=torch.int32)
        sb.sageattn_varlen(q[0].transpose(0, 1), q[0].transpose(0, 1), q[0].transpose(0, 1), cu, cu, 64, 64)
    with pytest.raises(RuntimeError):
        sb.per_warp_int8(q, q)
    with pytest.raises(RuntimeError):
        sb.per_channel_fp8(q)
    with pytest.raises(ValueError):
        sb.sageattn_qk_int8_pv_fp8_cuda(q.cuda() if torch.cuda.is_available() else q, q, q, pv_accum_dtype="bogus") \
            if torch.cuda.is_available() else (_ for _ in ()).throw(ValueError())


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sageattention_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("oracle/_ref", ""), f"{fn} references the oracle"
