// sage_attn_i8_f8: INT8 QK^T + online softmax + e4m3 P.V with per-channel V scales (include/sage_b200.h), the
// instantiations of sage_attn_fwd_kernel<D, /*kFp8=*/true, kMask> (attn_kernel_sm100.cuh).
// Replaces qk_int8_sv_f8_accum_f32_fuse_v_scale_attn and its siblings
// (/root/reference/csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh, qk_int_sv_f8_cuda_sm90.cu).
#include "attn_kernel_sm100.cuh"

#ifdef SAGE_TRACE
extern "C" int sage_debug_trace(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, sage::g_trace, sizeof(sage::g_trace));
}
#endif

template <int D>
static int launch_f8(const sage_attn_args* a, cudaStream_t s) {
  if (a->mask != nullptr) return sage_attn::launch_attn<D, true, true>(a, s);
  return sage_attn::launch_attn<D, true>(a, s);
}

extern "C" int sage_attn_i8_f8(const sage_attn_args* a, void* stream) {
  int rc = sage::check_attn_args(a);
  if (rc) return rc;
  SAGE_REQUIRE(a->cu_seqlens_q == nullptr, "varlen is an f16-P.V path (sage_attn_i8_f16)");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? launch_f8<128>(a, s) : launch_f8<64>(a, s);
}
