// sage_sageattn_fwd: the whole operator (quantisation prologue + attention) behind ONE C-ABI call with ONE caller-provided
// workspace, for the default paths of
//   sageattn / sageattn_qk_int8_pv_fp8_cuda   (/root/reference/sageattention/core.py:656-905: smooth_k, no smooth_v)
//   sageattn_qk_int8_pv_fp16_cuda             (/root/reference/sageattention/core.py:480-653: smooth_k, no smooth_v)
// The reference's Python makes 6-8 allocations and 3-5 extension calls per operator call; the binding of round 1 made
// 6-8 allocations and 2 calls (about 75 us of host time, more than the kernels take at N <= 2K).  Here the host side is
// two allocations (workspace, output) and one foreign call.  Up to kFlatPrologueRows kv rows it launches exactly what the
// two-call path launches (it calls the same entry points), so results are bit-identical to it; above, the prologue is
// issued as flat launches (below), whose K mean may differ from the cluster kernel's in the last bit of a channel.
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/sage_b200.h"
#include "status.h"

namespace {

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// Above this many kv rows the prologue is issued as flat launches (K mean, Q quantiser, K quantiser, V range + conversion)
// instead of the one-launch cluster prologue: a cluster of 8 CTAs owns a whole (b, h) slice, which is the right shape for
// many short slices but leaves most SMs idle on a few long ones (a sequence-parallel DiT shard: 5 heads x 75 600 rows =
// 40 CTAs per tensor; measured 352 -> 113 us for Q/K there, 805 -> 646 us with all 40 heads on one GPU), and beyond 64 K
// rows the slices in flight no longer fit L2, so the second read of K comes from HBM either way.  The choice depends on
// the sequence length ONLY, so a head-sharded run takes the same path - and gets the same bits - as the one-GPU run.
constexpr int kFlatPrologueRows = 65536;

struct Plan {
  size_t q8, k8, qs, ks, km, v, vs, stat, zmean, total;   // byte offsets into the workspace
  int q_per_bh, k_per_bh, npad;
  bool flat;
};

Plan make_plan(const sage_fwd_args* a) {
  Plan p{};
  const size_t B = a->B, Hq = a->Hq, Hkv = a->Hkv, Nq = a->Nq, Nk = a->Nk, D = a->D;
  p.q_per_bh = (a->Nq + 127) / 128 * (128 / a->warp_q) * (a->per_thread ? 8 : 1);
  p.k_per_bh = (a->Nk + 63) / 64 * (a->per_thread ? 4 : 1);
  // the reference zero-pads V to a multiple of 128 tokens before the fp8 conversion (core.py:822-828)
  p.npad = (a->Nk + 127) / 128 * 128;
  size_t off = 0;
  p.q8 = off; off = align256(off + B * Hq * Nq * D);
  p.k8 = off; off = align256(off + B * Hkv * Nk * D);
  p.qs = off; off = align256(off + B * Hq * (size_t)p.q_per_bh * 4);
  p.ks = off; off = align256(off + B * Hkv * (size_t)p.k_per_bh * 4);
  p.km = off; off = align256(off + B * Hkv * D * 2);
  p.v = off;
  const size_t d_og = a->d_og > 0 ? a->d_og : a->D;
  if (a->pv_fp8) off = align256(off + B * Hkv * D * (size_t)p.npad);
  else if (a->dtype == SAGE_DTYPE_BF16) off = align256(off + B * Hkv * Nk * d_og * 2);   // the f16 copy of a bf16 V
  p.vs = off;
  if (a->pv_fp8) off = align256(off + B * Hkv * D * 4);
  p.flat = a->Nk > kFlatPrologueRows && d_og == (size_t)a->D;
  p.stat = p.zmean = off;
  if (p.flat) {
    size_t need = sage_seq_mean_workspace_bytes(a->B, a->Hkv, a->Nk, a->D);
    if (a->pv_fp8) {
      const size_t nv = sage_per_channel_fp8_workspace_bytes(a->B, a->Hkv, a->Nk, a->D);
      if (nv > need) need = nv;
    }
    off = align256(off + need);
    p.zmean = off;
    if (!a->pv_fp8 && a->dtype == SAGE_DTYPE_BF16) off = align256(off + B * Hkv * D * 2);
  }
  p.total = off;
  return p;
}

// contiguous tensor of the caller's layout: [B,H,N,W] (HND) or [B,N,H,W] (NHD); strides in elements
sage_tensor4 make_t4(void* ptr, int layout_hnd, int64_t H, int64_t N, int64_t W) {
  sage_tensor4 t;
  t.ptr = ptr;
  if (layout_hnd) { t.stride_b = H * N * W; t.stride_h = N * W; t.stride_n = W; }
  else { t.stride_b = N * H * W; t.stride_h = W; t.stride_n = H * W; }
  return t;
}

int check(const sage_fwd_args* a) {
  SAGE_REQUIRE(a != nullptr, "null args");
  SAGE_REQUIRE(a->D == 64 || a->D == 128, "head_dim must be 64 or 128 after padding (got %d)", a->D);
  SAGE_REQUIRE(a->B > 0 && a->Hq > 0 && a->Hkv > 0 && a->Nq > 0 && a->Nk > 0, "empty problem");
  SAGE_REQUIRE(a->warp_q == 16 || a->warp_q == 32, "WARPQ must be 16 or 32 (got %d)", a->warp_q);
  SAGE_REQUIRE(a->dtype == SAGE_DTYPE_F16 || a->dtype == SAGE_DTYPE_BF16, "dtype must be f16 or bf16");
  SAGE_REQUIRE(a->q.ptr && a->k.ptr && a->v.ptr && a->o.ptr, "null tensor pointer");
  return 0;
}

}  // namespace

extern "C" size_t sage_sageattn_fwd_workspace_bytes(const sage_fwd_args* a) {
  if (check(a)) return 0;
  return make_plan(a).total;
}

extern "C" int sage_sageattn_fwd(const sage_fwd_args* a, void* stream) {
  int rc = check(a);
  if (rc) return rc;
  const Plan p = make_plan(a);
  SAGE_REQUIRE(a->workspace != nullptr && (reinterpret_cast<uintptr_t>(a->workspace) & 255) == 0,
               "workspace must be 256-byte aligned");
  SAGE_REQUIRE(a->workspace_bytes >= p.total, "workspace too small: %zu bytes given, %zu needed", a->workspace_bytes, p.total);
  uint8_t* ws = reinterpret_cast<uint8_t*>(a->workspace);
  const int d_og = a->d_og > 0 ? a->d_og : a->D;
  const int hnd = a->layout_hnd;
  sage_tensor4 q8 = make_t4(ws + p.q8, hnd, a->Hq, a->Nq, a->D);
  sage_tensor4 k8 = make_t4(ws + p.k8, hnd, a->Hkv, a->Nk, a->D);
  float* q_scale = reinterpret_cast<float*>(ws + p.qs);
  float* k_scale = reinterpret_cast<float*>(ws + p.ks);
  void* km = a->km_out ? a->km_out : static_cast<void*>(ws + p.km);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a->per_thread && p.q_per_bh != (a->Nq + a->warp_q - 1) / a->warp_q) {
    // per_warp: groups past the last row are never written by the quantiser (the two-call path hands it zeros)
    SAGE_CUDA(cudaMemsetAsync(q_scale, 0, (size_t)a->B * a->Hq * p.q_per_bh * 4, st));
  }

  sage_attn_args at{};
  at.q8 = q8; at.k8 = k8; at.o = a->o;
  at.q_scale = q_scale; at.k_scale = k_scale;
  at.lse = a->lse;
  at.B = a->B; at.Hq = a->Hq; at.Hkv = a->Hkv; at.Nq = a->Nq; at.Nk = a->Nk; at.D = a->D;
  at.gran = a->per_thread ? SAGE_GRAN_PER_THREAD : SAGE_GRAN_PER_WARP;
  at.warp_q = a->warp_q;
  at.is_causal = a->is_causal;
  at.out_dtype = a->dtype;
  at.sm_scale = a->sm_scale;
  at.d_og = d_og != a->D ? d_og : 0;
  at.max_ctas = a->max_ctas;

  if (p.flat) {
    // ---- long sequences: flat launches (see kFlatPrologueRows)
    void* stat = ws + p.stat;
    rc = sage_seq_mean(a->k, km, a->B, a->Hkv, a->Nk, a->D, a->dtype, stat, stream);
    if (rc) return rc;
    if (a->per_thread) {
      rc = sage_quant_per_thread_int8(a->q, nullptr, q8, q_scale, a->B, a->Hq, a->Nq, a->D, a->dtype, 0, a->warp_q, stream);
      if (rc) return rc;
      rc = sage_quant_per_thread_int8(a->k, km, k8, k_scale, a->B, a->Hkv, a->Nk, a->D, a->dtype, 1, 64, stream);
    } else {
      rc = sage_quant_int8(a->q, nullptr, q8, q_scale, a->B, a->Hq, a->Nq, a->D, a->dtype, a->warp_q, p.q_per_bh, 0, 0.f, stream);
      if (rc) return rc;
      rc = sage_quant_int8(a->k, km, k8, k_scale, a->B, a->Hkv, a->Nk, a->D, a->dtype, 64, p.k_per_bh, 0, 0.f, stream);
    }
    if (rc) return rc;
    if (a->pv_fp8) {
      uint8_t* v8 = ws + p.v;
      float* v_scale = reinterpret_cast<float*>(ws + p.vs);
      const int64_t np = p.npad;
      const int64_t sb = (int64_t)a->Hkv * a->D * np;
      const int64_t sh = hnd ? a->D * np : np;
      const int64_t sd = hnd ? np : (int64_t)a->Hkv * np;
      rc = sage_per_channel_fp8_flat(a->v, v8, sb, sh, sd, v_scale, a->B, a->Hkv, a->Nk, p.npad, a->D, a->dtype,
                                     a->v_scale_max, stat, stream);
      if (rc) return rc;
      at.v.ptr = v8; at.v.stride_b = sb; at.v.stride_h = sh; at.v.stride_n = sd;
      at.v_scale = v_scale;
      at.npad_v = p.npad;
    } else if (a->dtype == SAGE_DTYPE_BF16) {
      // the f16 copy of a bf16 V: half(v - 0) is the exact cast (the packed subtraction of zero is the identity)
      sage_tensor4 v16 = make_t4(ws + p.v, hnd, a->Hkv, a->Nk, d_og);
      void* zero = ws + p.zmean;
      SAGE_CUDA(cudaMemsetAsync(zero, 0, (size_t)a->B * a->Hkv * a->D * 2, st));
      rc = sage_sub_mean(a->v, zero, v16, a->B, a->Hkv, a->Nk, a->D, a->dtype, stream);
      if (rc) return rc;
      at.v = v16;
    } else {
      at.v = a->v;
    }
  } else if (a->pv_fp8) {
    // v_fp8: [B,Hkv,D,Npad] (HND) or [B,D,Hkv,Npad] (NHD), as the two-call path allocates it
    uint8_t* v8 = ws + p.v;
    float* v_scale = reinterpret_cast<float*>(ws + p.vs);
    const int64_t np = p.npad;
    const int64_t sb = (int64_t)a->Hkv * a->D * np;
    const int64_t sh = hnd ? a->D * np : np;
    const int64_t sd = hnd ? np : (int64_t)a->Hkv * np;
    rc = sage_prepass_qkv_fp8(a->q, a->k, q8, k8, q_scale, k_scale, km, a->v, v8, sb, sh, sd, v_scale, a->B, a->Hq, a->Hkv,
                              a->Nq, a->Nk, p.npad, a->D, a->dtype, a->per_thread, a->warp_q, p.q_per_bh, p.k_per_bh,
                              a->v_scale_max, d_og, stream);
    if (rc) return rc;
    at.v.ptr = v8; at.v.stride_b = sb; at.v.stride_h = sh; at.v.stride_n = sd;
    at.v_scale = v_scale;
    at.npad_v = p.npad;
  } else if (a->dtype == SAGE_DTYPE_BF16) {
    sage_tensor4 v16 = make_t4(ws + p.v, hnd, a->Hkv, a->Nk, d_og);
    rc = sage_prepass_qk_int8_v16(a->q, a->k, q8, k8, q_scale, k_scale, km, a->v, v16, a->B, a->Hq, a->Hkv, a->Nq, a->Nk,
                                  a->D, a->dtype, a->per_thread, a->warp_q, p.q_per_bh, p.k_per_bh, d_og, stream);
    if (rc) return rc;
    at.v = v16;
  } else {
    rc = sage_prepass_qk_int8(a->q, a->k, q8, k8, q_scale, k_scale, km, a->B, a->Hq, a->Hkv, a->Nq, a->Nk, a->D, a->dtype,
                              a->per_thread, a->warp_q, p.q_per_bh, p.k_per_bh, d_og, stream);
    if (rc) return rc;
    at.v = a->v;
  }
  if (a->ev_start) SAGE_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->ev_start), st));
  rc = a->pv_fp8 ? sage_attn_i8_f8(&at, stream) : sage_attn_i8_f16(&at, stream);
  if (rc) return rc;
  if (a->ev_stop) SAGE_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->ev_stop), st));
  return 0;
}
