// sage_attn_i8_f16: INT8 QK^T + online softmax + f16 P.V (include/sage_b200.h), the instantiations of
// sage_attn_fwd_kernel<D, /*kFp8=*/false, kMask, kVarlen> (attn_kernel_sm100.cuh).
// Replaces qk_int8_sv_f16_accum_f32_attn and its siblings (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:676-1500)
// and the Triton attention kernels behind attn_mask / varlen (/root/reference/sageattention/triton/attn_qk_int8_per_thread.py,
// attn_qk_int8_block_varlen.py).
#include "attn_kernel_sm100.cuh"

#ifdef SAGE_TRACE
extern "C" int sage_debug_trace(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, sage::g_trace, sizeof(sage::g_trace));
}
#endif

template <int D>
static int launch_f16(const sage_attn_args* a, cudaStream_t s) {
  if (a->cu_seqlens_q != nullptr) return sage_attn::launch_attn<D, false, false, true>(a, s);
  if (a->mask != nullptr) return sage_attn::launch_attn<D, false, true>(a, s);
  return sage_attn::launch_attn<D, false>(a, s);
}

extern "C" int sage_attn_i8_f16(const sage_attn_args* a, void* stream) {
  int rc = sage::check_attn_args(a);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? launch_f16<128>(a, s) : launch_f16<64>(a, s);
}
