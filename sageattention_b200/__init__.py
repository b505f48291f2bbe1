"""sageattention_b200 - Blackwell-native (sm_100a) SageAttention behind the reference's Python API.

Exports the same names as /root/reference/sageattention/__init__.py, as real distinct functions
(the fork aliases them all to one implementation, __init__.py:59-95).
"""
from .core import (sageattn, sageattn_qk_int8_pv_fp16_cuda, sageattn_qk_int8_pv_fp16_triton, sageattn_qk_int8_pv_fp8_cuda,
                   sageattn_varlen)
from .host import HostPipeline
from .parallel import shard_heads, ulysses_sageattn
from .quant import (per_block_int8, per_block_int8_varlen, per_channel_fp8, per_thread_int8, per_warp_int8, prepass_qk_int8, prepass_qkv_fp8, quant_q_int8, seq_mean,
                    smooth_quant_k_int8, sub_mean)

# the reference's sm90 entry point has the same contract as the fp8 one; on sm_100a both are the
# tcgen05 kernel (reference: core.py:908-1065)
sageattn_qk_int8_pv_fp8_cuda_sm90 = sageattn_qk_int8_pv_fp8_cuda

__all__ = [
    "sageattn",
    "sageattn_qk_int8_pv_fp16_cuda",
    "sageattn_qk_int8_pv_fp16_triton",
    "sageattn_qk_int8_pv_fp8_cuda",
    "sageattn_qk_int8_pv_fp8_cuda_sm90",
    "sageattn_varlen",
    "per_block_int8",
    "per_block_int8_varlen",
    "per_warp_int8",
    "per_thread_int8",
    "sub_mean",
    "per_channel_fp8",
    "HostPipeline",
    "seq_mean",
    "smooth_quant_k_int8",
    "quant_q_int8",
    "prepass_qk_int8",
    "prepass_qkv_fp8",
    "ulysses_sageattn",
    "shard_heads",
]
