// SageAttention forward for sm_100a (B200): INT8 QK^T + online softmax + FP8/FP16 P.V
//
// Replaces the reference's mma.sync kernels
//   qk_int_sv_f16_attn_kernel  (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671)
//   qk_int_sv_f8_attn_kernel   (/root/reference/csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh:44-707)
// with a from-scratch tcgen05 design (not a port):
//
//   * one CTA owns TWO 128-row Q tiles (256 query rows) of one (batch, head) and walks the KV
//     sequence in 64-row STEPS; K/V arrive by TMA in 128-row stages (two steps per stage) on a
//     3-deep mbarrier ring, 128B-swizzled;
//   * warp 8 is the TMA producer, warp 9 the single-thread MMA issuer:
//       S_t(j) = Q_t.K(j)^T   tcgen05.mma kind::i8, M=128 N=64, s32 accumulator in TMEM
//       O_t   += P_t(j).V(j)  kind::f8f6f4 (e4m3) or kind::f16, A operand (P) read from TMEM
//     S_t is DOUBLE-BUFFERED (two 64-column buffers per Q tile) so QK(j+2) is issued as soon as
//     PV(j) has been issued and the softmax never waits for the tensor pipe in steady state;
//   * warps 0-3 / 4-7 are two softmax warpgroups, one per Q tile, one thread per query row.
//     Per step: tcgen05.ld of S(j+1) is issued BEFORE the exp phase of S(j) (register double
//     buffer), integer per-k-scale-class max, dequant folded into one FFMA per element, ex2,
//     P packed to e4m3/f16 and written over the consumed S buffer with tcgen05.st.
//     The O accumulator is rescaled lazily: only when a row maximum grows by more than kTau
//     (exact softmax: numerator and denominator share the same reference maximum).
//
// TMEM map (512 columns):  S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D);
// buffer b of S_t = columns [64b, 64b+64); P_t(j) overwrites the first 16 (e4m3) / 32 (f16)
// columns of buffer j&1.
//
// Numerics contract (tolerance-checked against the reference kernels and fp32 SDPA):
//   s' = S * q_scale * k_scale * sm_scale * log2(e);  m = reference row max (>= true max - kTau);
//   p = ex2(s' - m [+ offset]);  d = sum of f32 p;  O = sum p.v (f32 in TMEM);
//   out = O / d [* v_scale] [+ v_mean];  lse = log2(d) + m [- offset]  (base 2, as the reference).
//   fp8: offset + kTau = 8.807 = the reference's exponent bias (attn_utils.cuh:30), so p <= 448.
//   Masked entries (causal kv > q, kv >= kv_len) contribute exactly 0.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/sage_b200.h"
#include "ptx_sm100.cuh"
#include "status.h"

namespace sage {

constexpr int kTileQ = 128;    // rows per Q tile (= TMEM lanes)
constexpr int kStepKV = 64;    // kv rows per MMA / softmax step
constexpr int kStageKV = 128;  // kv rows per TMA stage (two steps)
constexpr int kKStages = 3;
constexpr int kVStages = 3;
constexpr int kThreads = 384;  // 2 softmax warpgroups + 1 warpgroup holding the TMA warp and the MMA warp
constexpr int kSoftmaxRegs = 232;  // setmaxnreg: 2*128*232 + 128*40 <= 64K registers
constexpr int kIssueRegs = 40;
constexpr float kFp8Tau = 2.0f;                   // lazy-rescale slack (log2 units), fp8 P
constexpr float kFp8Offset = 8.807f - kFp8Tau;    // reference bias 8.807 (attn_utils.cuh:30) minus the slack
constexpr float kF16Tau = 8.0f;                   // f16 P: p <= 256
constexpr float kMagicF = 12582912.0f;            // 1.5 * 2^23
constexpr int kMagicI = 0x4B400000;
constexpr int kMaskedS = -4194304;  // below any reachable int8 dot product (|S| <= 2^21)

struct AttnKernelParams {
  const float* q_scale;
  const float* k_scale;
  const float* v_scale;  // may be null
  const float* v_mean;   // may be null
  void* o;
  float* lse;    // may be null
  int* debug_s;  // may be null: raw S of kv rows [0,128), [B,Hq,ceil256(Nq),128] s32
  int B, Hq, Hkv, Nq, Nk, Npad_v;
  long long o_stride_b, o_stride_h, o_stride_n;  // elements
  int q_scale_per_bh, k_scale_per_bh;            // entries per (b,h)
  int gran;                                      // 1 per_block(128/64) 2 per_warp 3 per_thread
  int warp_q;                                    // 16 or 32 rows per q-scale group (gran 2/3)
  int causal;
  int out_bf16;
  float sm_scale_log2;
};

template <int D, bool kFp8>
struct SmemLayout {
  static constexpr int kQBytes = kTileQ * D;                              // int8
  static constexpr int kKBytes = kStageKV * D;                            // int8
  static constexpr int kVBytes = kFp8 ? D * kStageKV : kStageKV * D * 2;  // e4m3 [D][128] | f16 [128][D]
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + 2 * kQBytes;
  static constexpr int kVOff = kKOff + kKStages * kKBytes;
  static constexpr int kBarOff = kVOff + kVStages * kVBytes;
  static constexpr int kNumBars = 1 + 2 * kKStages + 2 * kVStages + 4 + 4 + 2 + 2;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kTotal = kTmemPtrOff + 16;
  static constexpr int kDynamic = kTotal + 1024;  // slack for 1024B alignment
};

// K-scale class of S column c inside a 64-row step (SURVEY 8a Q3 / quant_per_thread.py:75):
// 4 interleaved groups.  per_warp / per_block use the same table with the 4 entries equal.
__device__ __forceinline__ constexpr int kcls(int c) { return (c & 7) >> 1; }

__device__ __forceinline__ uint32_t pack_e4m3x4(float a0, float a1, float a2, float a3) {
  // byte0 = a0, byte1 = a1, byte2 = a2, byte3 = a3
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(lo) : "f"(a1), "f"(a0));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(hi) : "f"(a3), "f"(a2));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
__device__ __forceinline__ uint32_t pack_f16x2(float a0, float a1) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(a1), "f"(a0));
  return r;
}

// ------------------------------------------------------------------------------------------
// Row maximum of one 64-column step (dequantised, log2 domain).  Integer max per k-scale class,
// then 4 multiplies.  kMasked: columns > lim are replaced by kMaskedS first.
template <bool kMasked>
__device__ __forceinline__ float step_max(uint32_t (&s)[64], const float (&sc)[4], int lim) {
  if constexpr (kMasked) {
#pragma unroll
    for (int c = 0; c < 64; ++c)
      if (c > lim) s[c] = (uint32_t)kMaskedS;
  }
  int mx[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) mx[k] = kMaskedS;
#pragma unroll
  for (int c = 0; c < 64; ++c) mx[kcls(c)] = max(mx[kcls(c)], (int)s[c]);
  float mrow = -1e30f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float cand = (float)mx[k] * sc[k];
    if constexpr (kMasked) cand = (mx[k] == kMaskedS) ? -1e30f : cand;  // an all-masked class must not win
    mrow = fmaxf(mrow, cand);
  }
  return mrow;
}

// p = ex2(S*sc - m_ref + offset) for the 64 columns of one step, packed into w[] for the PV MMA's
// A operand; returns the f32 row sum.  fp8: V^T is stored with the reference's 16-token
// permutation out[j] = in[{0,1,8,9,2,3,10,11,4,5,12,13,6,7,14,15}[j]]
// (/root/reference/csrc/fused/fused.cu:284-291), so P's K index is permuted the same way:
// word q of a 16-group = tokens (2q, 2q+1, 2q+8, 2q+9).
template <bool kFp8, bool kMasked>
__device__ __forceinline__ float step_exp(const uint32_t (&s)[64], const float (&sc)[4], float m_ref,
                                          int lim, uint32_t (&w)[kFp8 ? 16 : 32]) {
  float bias[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) bias[k] = fmaf(-kMagicF, sc[k], (kFp8 ? kFp8Offset : 0.f) - m_ref);
  float dsum0 = 0.f, dsum1 = 0.f;
  auto pexp = [&](int c) -> float {
    const float x = __int_as_float((int)s[c] + kMagicI);
    float p = ex2_approx(fmaf(x, sc[kcls(c)], bias[kcls(c)]));
    if constexpr (kMasked) p = (c > lim) ? 0.f : p;
    return p;
  };
  if constexpr (kFp8) {
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int t = g * 16 + 2 * q;
        const float p0 = pexp(t), p1 = pexp(t + 1), p2 = pexp(t + 8), p3 = pexp(t + 9);
        dsum0 += p0 + p2;
        dsum1 += p1 + p3;
        w[g * 4 + q] = pack_e4m3x4(p0, p1, p2, p3);
      }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float p0 = pexp(2 * i), p1 = pexp(2 * i + 1);
      dsum0 += p0;
      dsum1 += p1;
      w[i] = pack_f16x2(p0, p1);
    }
  }
  return dsum0 + dsum1;
}

template <int D, bool kFp8>
struct SoftmaxCtx {
  const AttnKernelParams& p;
  int t, q_row, q_tile_start, n_iter;
  uint32_t tS, tO;
  float qs;
  const float* ksp;
  uint64_t *s_full, *p_full, *pv_done;
  float m_ref, d_run;
};

template <int D, bool kFp8>
__device__ __forceinline__ void load_k_scales(const SoftmaxCtx<D, kFp8>& c, int j, float (&sc)[4]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int idx = (c.p.gran == 3) ? j * 4 + k : j;  // per_warp / per_block: one scale per 64 kv rows
    idx = min(idx, c.p.k_scale_per_bh - 1);
    sc[k] = __ldg(c.ksp + idx) * c.qs;
  }
}

// One softmax step for one query row.  `cur` holds S(j) (already loaded, row max `mt` known);
// `nxt` receives S(j+1).  On exit mt / scc describe step j+1.
template <int D, bool kFp8>
__device__ __forceinline__ void softmax_step(SoftmaxCtx<D, kFp8>& c, int j, uint32_t (&cur)[64],
                                             uint32_t (&nxt)[64], float (&scc)[4], float& mt) {
  const AttnKernelParams& p = c.p;
  const bool has_next = (j + 1 < c.n_iter);
  float scn[4];
  if (has_next) load_k_scales(c, j + 1, scn);

  // ---- reference maximum, lazy O rescale (PV(j-1) must be complete before O is touched)
  if (j == 0) {
    c.m_ref = mt;
  } else {
    const bool grow = mt > c.m_ref + (kFp8 ? kFp8Tau : kF16Tau);
    if (__any_sync(0xffffffffu, grow)) {
      const float m_new = fmaxf(c.m_ref, mt);
      const float alpha = ex2_approx(c.m_ref - m_new);
      c.m_ref = m_new;
      c.d_run *= alpha;
      mbar_wait(&c.pv_done[c.t], (j - 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < D; c0 += 32) {
        uint32_t o[32];
        tmem_ld32(c.tO + c0, o);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
        tmem_st32(c.tO + c0, o);
      }
    }
  }

  // ---- prefetch S(j+1) into registers while the exp phase of S(j) runs
  if (has_next) {
    mbar_wait(&c.s_full[c.t * 2 + ((j + 1) & 1)], ((j + 1) >> 1) & 1);
    tc_fence_after();
    const uint32_t src = c.tS + (uint32_t)(((j + 1) & 1) * 64);
    uint32_t(&n0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&nxt[0]);
    uint32_t(&n1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&nxt[32]);
    tmem_ld32(src, n0);
    tmem_ld32(src + 32, n1);
  }

  // ---- exp / pack / store P(j)
  const int kv0 = j * kStepKV;
  int lim = p.Nk - 1 - kv0;
  if (p.causal) lim = min(lim, c.q_row - kv0);
  const bool masked = (kv0 + kStepKV > p.Nk) || (p.causal && kv0 + kStepKV - 1 > c.q_tile_start);
  uint32_t w[kFp8 ? 16 : 32];
  float dsum;
  if (masked) dsum = step_exp<kFp8, true>(cur, scc, c.m_ref, lim, w);
  else dsum = step_exp<kFp8, false>(cur, scc, c.m_ref, 63, w);
  c.d_run += dsum;
  const uint32_t tP = c.tS + (uint32_t)((j & 1) * 64);
  if constexpr (kFp8) tmem_st16(tP, w);
  else tmem_st32(tP, w);
  tc_wait_st();
  tc_fence_before();
  mbar_arrive(&c.p_full[c.t * 2 + (j & 1)]);

  // ---- row max of S(j+1)
  if (has_next) {
    tc_wait_ld();
    const int kv1 = kv0 + kStepKV;
    int lim1 = p.Nk - 1 - kv1;
    if (p.causal) lim1 = min(lim1, c.q_row - kv1);
    const bool masked1 = (kv1 + kStepKV > p.Nk) || (p.causal && kv1 + kStepKV - 1 > c.q_tile_start);
    if (masked1) mt = step_max<true>(nxt, scn, lim1);
    else mt = step_max<false>(nxt, scn, 63);
#pragma unroll
    for (int k = 0; k < 4; ++k) scc[k] = scn[k];
  }
}

// Softmax warpgroup body for Q tile `t` (128 threads, thread = query row).
template <int D, bool kFp8>
__device__ __forceinline__ void softmax_warpgroup(const AttnKernelParams& p, int t, int row,
                                                  int q_tile_start, int n_iter, int b, int h,
                                                  int hkv, uint32_t tmem_base, uint64_t* s_full,
                                                  uint64_t* p_full, uint64_t* pv_done,
                                                  uint64_t* o_full) {
  const int warp_in_wg = (row >> 5);
  const uint32_t lane_off = (uint32_t)(warp_in_wg * 32) << 16;
  SoftmaxCtx<D, kFp8> c{p};
  c.t = t;
  c.q_row = q_tile_start + row;
  c.q_tile_start = q_tile_start;
  c.n_iter = n_iter;
  c.tS = tmem_base + lane_off + (uint32_t)(t * 128);
  c.tO = tmem_base + lane_off + (uint32_t)(256 + t * D);
  c.s_full = s_full;
  c.p_full = p_full;
  c.pv_done = pv_done;
  c.m_ref = -1e30f;
  c.d_run = 0.f;
  const int q_row = c.q_row;

  // per-row q scale (folded with sm_scale * log2 e)
  {
    int idx;
    if (p.gran == 3) idx = (q_row / p.warp_q) * 8 + (q_row & 7);
    else if (p.gran == 2) idx = q_row / p.warp_q;
    else idx = q_row / 128;
    idx = min(idx, p.q_scale_per_bh - 1);
    c.qs = p.q_scale[(size_t)(b * p.Hq + h) * p.q_scale_per_bh + idx] * p.sm_scale_log2;
  }
  c.ksp = p.k_scale + (size_t)(b * p.Hkv + hkv) * p.k_scale_per_bh;

  uint32_t bufA[64], bufB[64];
  float scc[4], mt;
  // ---- prologue: S(0)
  load_k_scales(c, 0, scc);
  mbar_wait(&s_full[t * 2 + 0], 0);
  tc_fence_after();
  {
    uint32_t(&a0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&bufA[0]);
    uint32_t(&a1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&bufA[32]);
    tmem_ld32(c.tS, a0);
    tmem_ld32(c.tS + 32, a1);
    tc_wait_ld();
  }
  if (p.debug_s != nullptr) {
    const int nq_pad = (p.Nq + 255) / 256 * 256;
    int* dst = p.debug_s + ((size_t)(b * p.Hq + h) * nq_pad + q_row) * 128;
#pragma unroll
    for (int cc = 0; cc < 64; ++cc) dst[cc] = (int)bufA[cc];
  }
  {
    int lim0 = p.Nk - 1;
    if (p.causal) lim0 = min(lim0, q_row);
    const bool masked0 = (kStepKV > p.Nk) || (p.causal && kStepKV - 1 > q_tile_start);
    if (masked0) mt = step_max<true>(bufA, scc, lim0);
    else mt = step_max<false>(bufA, scc, 63);
  }
  for (int j = 0; j < n_iter; j += 2) {
    softmax_step<D, kFp8>(c, j, bufA, bufB, scc, mt);
    if (p.debug_s != nullptr && j == 0 && n_iter > 1) {
      const int nq_pad = (p.Nq + 255) / 256 * 256;
      int* dst = p.debug_s + ((size_t)(b * p.Hq + h) * nq_pad + q_row) * 128 + 64;
#pragma unroll
      for (int cc = 0; cc < 64; ++cc) dst[cc] = (int)bufB[cc];
    }
    if (j + 1 < n_iter) softmax_step<D, kFp8>(c, j + 1, bufB, bufA, scc, mt);
  }

  // ---- epilogue
  mbar_wait(&o_full[t], 0);
  tc_fence_after();
  const float inv_d = rcp_approx(c.d_run);
  const bool store = q_row < p.Nq;
  const float* vsp = p.v_scale ? p.v_scale + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  const float* vmp = p.v_mean ? p.v_mean + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  uint16_t* optr = reinterpret_cast<uint16_t*>(p.o) + (size_t)b * p.o_stride_b +
                   (size_t)h * p.o_stride_h + (size_t)q_row * p.o_stride_n;
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 32) {
    uint32_t o[32];
    tmem_ld32(c.tO + c0, o);
    tc_wait_ld();
    uint32_t packed[16];
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float x0 = __uint_as_float(o[i]) * inv_d;
      float x1 = __uint_as_float(o[i + 1]) * inv_d;
      if (vsp) {
        x0 *= __ldg(vsp + c0 + i);
        x1 *= __ldg(vsp + c0 + i + 1);
      }
      if (vmp) {
        x0 += __ldg(vmp + c0 + i);
        x1 += __ldg(vmp + c0 + i + 1);
      }
      if (p.out_bf16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(x0, x1);
        packed[i >> 1] = *reinterpret_cast<uint32_t*>(&v);
      } else {
        __half2 v = __floats2half2_rn(x0, x1);
        packed[i >> 1] = *reinterpret_cast<uint32_t*>(&v);
      }
    }
    if (store) {
      uint4* dst = reinterpret_cast<uint4*>(optr + c0);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        dst[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
    }
  }
  if (p.lse && store)
    p.lse[(size_t)(b * p.Hq + h) * p.Nq + q_row] =
        lg2_approx(c.d_run) + c.m_ref - (kFp8 ? kFp8Offset : 0.f);
}

// ------------------------------------------------------------------------------------------
template <int D, bool kFp8>
__global__ void __launch_bounds__(kThreads, 1)
sage_attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnKernelParams p) {
  using L = SmemLayout<D, kFp8>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* q_full = bars;                    // 1
  uint64_t* k_full = q_full + 1;              // kKStages
  uint64_t* k_empty = k_full + kKStages;      // kKStages
  uint64_t* v_full = k_empty + kKStages;      // kVStages
  uint64_t* v_empty = v_full + kVStages;      // kVStages
  uint64_t* s_full = v_empty + kVStages;      // 4: [tile][buffer]
  uint64_t* p_full = s_full + 4;              // 4: [tile][buffer]
  uint64_t* pv_done = p_full + 4;             // 2
  uint64_t* o_full = pv_done + 2;             // 2
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);

  const int warp = threadIdx.x >> 5;
  // heavy (late) q tiles first: helps the causal tail
  const int q_pair = (int)gridDim.x - 1 - (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hkv = h / (p.Hq / p.Hkv);
  const int q_start = q_pair * 2 * kTileQ;

  // kv steps per q tile (causal: stop at the diagonal; top-left aligned as the reference,
  // /root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:218-222)
  int n_it[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int qs = q_start + t * kTileQ;
    int kv_end = p.Nk;
    if (p.causal) kv_end = min(kv_end, qs + kTileQ);
    n_it[t] = (qs < p.Nq) ? (kv_end + kStepKV - 1) / kStepKV : 0;
  }
  const int n_max = max(n_it[0], n_it[1]);
  const int n_stages = (n_max + 1) / 2;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < kKStages; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < kVStages; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], kTileQ); }
    for (int t = 0; t < 2; ++t) { mbar_init(&pv_done[t], 1); mbar_init(&o_full[t], 1); }
    fence_barrier_init();
  }
  if (warp == 8) {
    if (elect_one()) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
    }
  }
  if (warp == 9) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp < 8) {
    // ======================= softmax warpgroups =======================
    reg_alloc<kSoftmaxRegs>();
    const int t = warp >> 2;
    const int row = threadIdx.x & 127;
    if (n_it[t] > 0)
      softmax_warpgroup<D, kFp8>(p, t, row, q_start + t * kTileQ, n_it[t], b, h, hkv, tmem_base,
                                 s_full, p_full, pv_done, o_full);
  } else if (warp == 8) {
    // ======================= TMA producer =======================
    reg_dealloc<kIssueRegs>();
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, 2 * L::kQBytes);
      tma_load_4d(smem + L::kQOff, &tmQ, q_full, 0, q_start, h, b);
      tma_load_4d(smem + L::kQOff + L::kQBytes, &tmQ, q_full, 0, q_start + kTileQ, h, b);
      for (int s = 0; s < n_stages; ++s) {
        const int ks = s % kKStages, vs = s % kVStages;
        mbar_wait(&k_empty[ks], ((s / kKStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&k_full[ks], L::kKBytes);
        tma_load_4d(smem + L::kKOff + ks * L::kKBytes, &tmK, &k_full[ks], 0, s * kStageKV, hkv, b);
        mbar_wait(&v_empty[vs], ((s / kVStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&v_full[vs], L::kVBytes);
        uint8_t* vdst = smem + L::kVOff + vs * L::kVBytes;
        if constexpr (kFp8) {
          // V^T e4m3 [D rows][128 kv] (kv contiguous)
          tma_load_4d(vdst, &tmV, &v_full[vs], s * kStageKV, 0, hkv, b);
        } else {
          // V f16 [128 kv][D]: one 64-wide (128 B) column block per load
#pragma unroll
          for (int cb = 0; cb < D / 64; ++cb)
            tma_load_4d(vdst + cb * (kStageKV * 128), &tmV, &v_full[vs], cb * 64, s * kStageKV, hkv, b);
        }
      }
    }
  } else if (warp > 9) {
    reg_dealloc<kIssueRegs>();  // idle warps of the issue warpgroup (setmaxnreg is warpgroup-wide)
  } else {
    // ======================= MMA issuer =======================
    reg_dealloc<kIssueRegs>();
    if (elect_one()) {
      constexpr uint32_t idesc_qk = make_idesc(CFMT_S32, FMT_S8, FMT_S8, 0, 0, kTileQ, kStepKV);
      constexpr uint32_t idesc_pv =
          kFp8 ? make_idesc(CFMT_F32, FMT_E4M3, FMT_E4M3, 0, 0, kTileQ, D)
               : make_idesc(CFMT_F32, FMT_F16, FMT_F16, 0, /*B MN-major*/ 1, kTileQ, D);
      // Q/K rows are D bytes: one 128B (D=128) or 64B (D=64) swizzle span; 8-row groups are dense
      constexpr uint64_t qk_swz = (D == 128) ? SWZ_128B : SWZ_64B;
      constexpr uint32_t qk_sbo = 8 * D;
      const uint32_t sQ = smem_u32(smem + L::kQOff);
      const uint32_t sK = smem_u32(smem + L::kKOff);
      const uint32_t sV = smem_u32(smem + L::kVOff);

      // S_t(q) -> buffer q&1 ; K rows (q&1)*64.. of stage (q>>1)
      auto issue_qk = [&](int t, int q) {
        const uint32_t a0 = sQ + t * L::kQBytes;
        const uint32_t b0 = sK + ((q >> 1) % kKStages) * L::kKBytes + (q & 1) * (kStepKV * D);
        const uint32_t d_tmem = tmem_base + t * 128 + (q & 1) * 64;
#pragma unroll
        for (int kk = 0; kk < D / 32; ++kk) {
          const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, qk_sbo, qk_swz);
          const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, qk_sbo, qk_swz);
          mma_i8_ss(d_tmem, ad, bd, idesc_qk, kk > 0);
        }
      };
      auto issue_pv = [&](int t, int j) {
        const uint32_t b0 = sV + ((j >> 1) % kVStages) * L::kVBytes;
        const uint32_t d_tmem = tmem_base + 256 + t * D;
        const uint32_t a_tmem = tmem_base + t * 128 + (j & 1) * 64;
        if constexpr (kFp8) {
#pragma unroll
          for (int kk = 0; kk < kStepKV / 32; ++kk) {
            const uint64_t bd = make_smem_desc(b0 + (j & 1) * 64 + kk * 32, 16, 1024, SWZ_128B);
            mma_f8_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, (j > 0) || (kk > 0));
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < kStepKV / 16; ++kk) {
            // MN-major B: 16 kv rows x 128 B per MMA = 2 KB step; 64-wide d blocks 16 KB apart
            const uint64_t bd = make_smem_desc(b0 + (j & 1) * (kStepKV * 128) + kk * 2048, kStageKV * 128,
                                               1024, SWZ_128B);
            mma_f16_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, (j > 0) || (kk > 0));
          }
        }
      };
      // a K stage is free once the QKs of its last step have been issued for both tiles
      auto release_k_after = [&](int q) {
        if ((q & 1) == 1 || q == n_max - 1) tc_commit(&k_empty[(q >> 1) % kKStages]);
      };

      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      for (int q = 0; q < 2 && q < n_max; ++q) {
        for (int t = 0; t < 2; ++t)
          if (q < n_it[t]) { issue_qk(t, q); tc_commit(&s_full[t * 2 + q]); }
        release_k_after(q);
      }

      for (int j = 0; j < n_max; ++j) {
        if ((j & 1) == 0) mbar_wait(&v_full[(j >> 1) % kVStages], ((j >> 1) / kVStages) & 1);
        const int q = j + 2;
        if (q < n_max && (q & 1) == 0) mbar_wait(&k_full[(q >> 1) % kKStages], ((q >> 1) / kKStages) & 1);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (j < n_it[t]) {
            mbar_wait(&p_full[t * 2 + (j & 1)], (j >> 1) & 1);
            tc_fence_after();
            issue_pv(t, j);
            tc_commit(&pv_done[t]);
            if (j == n_it[t] - 1) tc_commit(&o_full[t]);
          }
          if (q < n_it[t]) { issue_qk(t, q); tc_commit(&s_full[t * 2 + (q & 1)]); }
        }
        if ((j & 1) == 1 || j == n_max - 1) tc_commit(&v_empty[(j >> 1) % kVStages]);
        if (q < n_max) release_k_after(q);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace sage

// ==========================================================================================
// Host side: TMA descriptors + launch
// ==========================================================================================
namespace sage {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// 4-D map: dims (inner, n, h, b); strides in BYTES for dims 1..3.
static int make_map4(CUtensorMap* m, CUtensorMapDataType dt, void* ptr, uint64_t d0, uint64_t d1,
                     uint64_t d2, uint64_t d3, uint64_t s1, uint64_t s2, uint64_t s3, uint32_t b0,
                     uint32_t b1, CUtensorMapSwizzle swz, const char* name) {
  EncodeTiledFn enc = get_encode_fn();
  SAGE_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
  SAGE_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "%s: base pointer must be 16-byte aligned", name);
  SAGE_REQUIRE(s1 % 16 == 0 && s2 % 16 == 0 && s3 % 16 == 0,
               "%s: strides must be multiples of 16 bytes (got %llu %llu %llu)", name,
               (unsigned long long)s1, (unsigned long long)s2, (unsigned long long)s3);
  cuuint64_t dims[4] = {d0, d1, d2, d3};
  cuuint64_t strides[3] = {s1, s2, s3};
  cuuint32_t box[4] = {b0, b1, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, dt, 4, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SAGE_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(%s) failed with CUresult %d", name, (int)r);
  return 0;
}

template <int D, bool kFp8>
static int launch_attn(const sage_attn_args* a, cudaStream_t stream) {
  using L = SmemLayout<D, kFp8>;
  CUtensorMap tmQ, tmK, tmV;
  const CUtensorMapSwizzle qk_swz = (D == 128) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  int rc;
  rc = make_map4(&tmQ, CU_TENSOR_MAP_DATA_TYPE_UINT8, a->q8.ptr, D, a->Nq, a->Hq, a->B,
                 a->q8.stride_n, a->q8.stride_h, a->q8.stride_b, D, kTileQ, qk_swz, "q_int8");
  if (rc) return rc;
  rc = make_map4(&tmK, CU_TENSOR_MAP_DATA_TYPE_UINT8, a->k8.ptr, D, a->Nk, a->Hkv, a->B,
                 a->k8.stride_n, a->k8.stride_h, a->k8.stride_b, D, kStageKV, qk_swz, "k_int8");
  if (rc) return rc;
  if constexpr (kFp8) {
    // v: e4m3 [B,Hkv,D,Npad] rows; tensor4 strides are reused as (b, h, d-row)
    SAGE_REQUIRE(a->npad_v >= a->Nk && a->npad_v % 64 == 0, "v_fp8 padded length %d invalid for kv_len %d", a->npad_v, a->Nk);
    rc = make_map4(&tmV, CU_TENSOR_MAP_DATA_TYPE_UINT8, a->v.ptr, a->npad_v, D, a->Hkv, a->B,
                   a->v.stride_n, a->v.stride_h, a->v.stride_b, kStageKV, D,
                   CU_TENSOR_MAP_SWIZZLE_128B, "v_fp8");
  } else {
    rc = make_map4(&tmV, CU_TENSOR_MAP_DATA_TYPE_UINT16, a->v.ptr, D, a->Nk, a->Hkv, a->B,
                   a->v.stride_n * 2, a->v.stride_h * 2, a->v.stride_b * 2, 64, kStageKV,
                   CU_TENSOR_MAP_SWIZZLE_128B, "v_f16");
  }
  if (rc) return rc;

  AttnKernelParams p;
  p.q_scale = a->q_scale;
  p.k_scale = a->k_scale;
  p.v_scale = a->v_scale;
  p.v_mean = a->v_mean;
  p.o = a->o.ptr;
  p.lse = a->lse;
  p.debug_s = reinterpret_cast<int*>(a->debug_s);
  p.B = a->B; p.Hq = a->Hq; p.Hkv = a->Hkv; p.Nq = a->Nq; p.Nk = a->Nk; p.Npad_v = a->npad_v;
  p.o_stride_b = a->o.stride_b; p.o_stride_h = a->o.stride_h; p.o_stride_n = a->o.stride_n;
  p.gran = a->gran;
  p.warp_q = a->warp_q;
  const int nqb = (a->Nq + 127) / 128, nkb = (a->Nk + 63) / 64;
  if (a->gran == SAGE_GRAN_PER_THREAD) {
    p.q_scale_per_bh = nqb * (128 / a->warp_q) * 8;
    p.k_scale_per_bh = nkb * 4;
  } else if (a->gran == SAGE_GRAN_PER_WARP) {
    p.q_scale_per_bh = nqb * (128 / a->warp_q);
    p.k_scale_per_bh = nkb;
  } else {
    p.q_scale_per_bh = nqb;
    p.k_scale_per_bh = nkb;
  }
  p.causal = a->is_causal;
  p.out_bf16 = (a->out_dtype == SAGE_DTYPE_BF16);
  p.sm_scale_log2 = a->sm_scale * 1.44269504088896340736f;

  auto kern = sage_attn_fwd_kernel<D, kFp8>;
  static bool attr_set = false;  // per template instantiation
  if (!attr_set) {
    SAGE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kDynamic));
    attr_set = true;
  }
  dim3 grid((a->Nq + 2 * kTileQ - 1) / (2 * kTileQ), a->Hq, a->B);
  kern<<<grid, kThreads, L::kDynamic, stream>>>(tmQ, tmK, tmV, p);
  SAGE_CUDA(cudaGetLastError());
  return 0;
}

static int check_attn_args(const sage_attn_args* a) {
  SAGE_REQUIRE(a != nullptr, "null args");
  SAGE_REQUIRE(a->D == 64 || a->D == 128, "head_dim must be 64 or 128 after padding (got %d)", a->D);
  SAGE_REQUIRE(a->B > 0 && a->Hq > 0 && a->Hkv > 0 && a->Nq > 0 && a->Nk > 0, "empty problem");
  SAGE_REQUIRE(a->Hq % a->Hkv == 0, "num_qo_heads (%d) must be divisible by num_kv_heads (%d)", a->Hq, a->Hkv);
  SAGE_REQUIRE(a->gran >= 1 && a->gran <= 3, "unsupported qk_quant_gran %d", a->gran);
  SAGE_REQUIRE(a->gran == 1 || a->warp_q == 16 || a->warp_q == 32, "warp_q must be 16 or 32");
  SAGE_REQUIRE(a->q8.ptr && a->k8.ptr && a->v.ptr && a->o.ptr && a->q_scale && a->k_scale, "null tensor pointer");
  SAGE_REQUIRE(a->o.stride_n % 8 == 0 && a->o.stride_h % 8 == 0 && a->o.stride_b % 8 == 0 &&
               (reinterpret_cast<uintptr_t>(a->o.ptr) & 15) == 0, "output rows must be 16-byte aligned");
  SAGE_REQUIRE(a->B <= 65535 && a->Hq <= 65535, "batch / heads exceed grid limits");
  return 0;
}

}  // namespace sage

extern "C" int sage_attn_i8_f8(const sage_attn_args* a, void* stream) {
  int rc = sage::check_attn_args(a);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? sage::launch_attn<128, true>(a, s) : sage::launch_attn<64, true>(a, s);
}

extern "C" int sage_attn_i8_f16(const sage_attn_args* a, void* stream) {
  int rc = sage::check_attn_args(a);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? sage::launch_attn<128, false>(a, s) : sage::launch_attn<64, false>(a, s);
}
