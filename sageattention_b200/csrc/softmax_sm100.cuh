// Per-row softmax arithmetic of the sm_100a attention kernel (shared by attn_sm100.cu and the
// stand-alone softmax microbenchmark tools/softmax_bench.cu).
#pragma once
#include <cstdint>
#include "ptx_sm100.cuh"

namespace sage {

constexpr float kFp8Tau = 2.0f;
constexpr float kF16Tau = 8.0f;
constexpr float kFp8Offset = 8.807f - kFp8Tau;  // log2(448) rounded down, as the reference (attn_utils.cuh:30)
constexpr float kMagicF = 12582912.0f;  // 1.5 * 2^23
constexpr int kMagicI = 0x4B400000;
constexpr int kMaskedS = -4194304;  // below any reachable int8 dot product (|S| <= 2^21)

// K-scale class of S column c inside a 128-wide KV tile (see SURVEY 8a Q3 / quant_per_thread.py:75):
// two 64-row blocks, 4 interleaved groups each.  per_warp / per_block use the same table with the
// 4 entries of a block set equal.
__device__ __forceinline__ constexpr int kcls(int c) { return (c >> 6) * 4 + ((c & 7) >> 1); }

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
// One softmax step for one query row: s[] holds the 128 int32 logits of this KV tile on entry and
// is reused for p.  `lim` = last visible column (127 when nothing is masked).
// Returns packed P words in w[] (32 words e4m3 / 64 words f16).
template <bool kFp8, bool kMasked>
__device__ __forceinline__ void softmax_row(uint32_t (&s)[128], const float (&sc)[8], int lim,
                                            float& m_run, float& d_run, float& alpha,
                                            uint32_t (&w)[kFp8 ? 32 : 64]) {
  if constexpr (kMasked) {
#pragma unroll
    for (int c = 0; c < 128; ++c)
      if (c > lim) s[c] = (uint32_t)kMaskedS;
  }
  // integer max per k-scale class, then 8 dequantised candidates
  int mx[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) mx[k] = kMaskedS;
#pragma unroll
  for (int c = 0; c < 128; ++c) mx[kcls(c)] = max(mx[kcls(c)], (int)s[c]);
  float mrow = -1e30f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    // a class whose columns are all masked must not win the max
    float cand = (float)mx[k] * sc[k];
    if constexpr (kMasked) cand = (mx[k] == kMaskedS) ? -1e30f : cand;
    mrow = fmaxf(mrow, cand);
  }
  const bool grow = mrow > m_run + (kFp8 ? kFp8Tau : kF16Tau);
  float m_new = m_run;
  alpha = 1.0f;
  if (__any_sync(0xffffffffu, grow)) {
    m_new = fmaxf(m_run, mrow);
    alpha = ex2_approx(m_run - m_new);  // 0 on the first tile (m_run = -1e30)
  }
  m_run = m_new;
  float bias[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) bias[k] = fmaf(-kMagicF, sc[k], (kFp8 ? kFp8Offset : 0.f) - m_new);
  float dsum0 = 0.f, dsum1 = 0.f;
#pragma unroll
  for (int c = 0; c < 128; c += 2) {
    float x0 = __int_as_float((int)s[c] + kMagicI);
    float x1 = __int_as_float((int)s[c + 1] + kMagicI);
    float p0 = ex2_approx(fmaf(x0, sc[kcls(c)], bias[kcls(c)]));
    float p1 = ex2_approx(fmaf(x1, sc[kcls(c + 1)], bias[kcls(c + 1)]));
    if constexpr (kMasked) {
      p0 = (c > lim) ? 0.f : p0;
      p1 = (c + 1 > lim) ? 0.f : p1;
    }
    dsum0 += p0;
    dsum1 += p1;
    s[c] = __float_as_uint(p0);
    s[c + 1] = __float_as_uint(p1);
  }
  d_run = fmaf(d_run, alpha, dsum0 + dsum1);
  if constexpr (kFp8) {
    // V^T is stored with the reference's 16-token permutation
    // out[j] = in[{0,1,8,9,2,3,10,11,4,5,12,13,6,7,14,15}[j]]  (/root/reference/csrc/fused/fused.cu:284-291),
    // so P's K index is permuted the same way: word q of a 16-group = tokens (2q, 2q+1, 2q+8, 2q+9).
#pragma unroll
    for (int g = 0; g < 8; ++g)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int t = g * 16 + 2 * q;
        w[g * 4 + q] = pack_e4m3x4(__uint_as_float(s[t]), __uint_as_float(s[t + 1]),
                                   __uint_as_float(s[t + 8]), __uint_as_float(s[t + 9]));
      }
  } else {
#pragma unroll
    for (int i = 0; i < 64; ++i)
      w[i] = pack_f16x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1]));
  }
}

}  // namespace sage
