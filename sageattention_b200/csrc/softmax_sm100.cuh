// Per-row softmax arithmetic of the sm_100a attention kernel (shared by attn_kernel_sm100.cuh and the
// stand-alone softmax microbenchmark tools/softmax_bench.cu).
#pragma once
#include <cstdint>
#include <cuda_fp16.h>
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
  // byte0 = a0, byte1 = a1, byte2 = a2, byte3 = a3.  The two halves are joined with a vector mov so
  // that ptxas uses F2FP's merge form (two instructions, no PRMT).
  uint32_t r;
  asm("{ .reg .b16 lo, hi;\n"
      "  cvt.rn.satfinite.e4m3x2.f32 lo, %2, %1;\n"
      "  cvt.rn.satfinite.e4m3x2.f32 hi, %4, %3;\n"
      "  mov.b32 %0, {lo, hi}; }\n"
      : "=r"(r)
      : "f"(a0), "f"(a1), "f"(a2), "f"(a3));
  return r;
}
__device__ __forceinline__ uint32_t pack_f16x2(float a0, float a1) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(a1), "f"(a0));
  return r;
}

// ------------------------------------------------------------------------------------------
// One softmax step for one query row, in three phases so the caller can schedule them (the exp
// phase is the only one that needs the MUFU pipe).  s[] holds the 128 int32 logits of this KV tile
// on entry and is reused for p.  `lim` = last visible column (127 when nothing is masked).

// Phase 1: (masking,) row max, lazy-rescale decision.  Out: m_run updated, alpha (1 when the
// accumulators keep their scale), bias[] for phase 2.
// kBiased: the logits arrive as S + kMagicI (the accumulator was pre-set to the magic constant, so
// the bits ARE the float 1.5*2^23 + S and no per-element int->float add is needed).
template <bool kFp8, bool kMasked, bool kBiased = false>
__device__ __forceinline__ void softmax_max(uint32_t (&s)[128], const float (&sc)[8], int lim,
                                            float& m_run, float& alpha, float (&bias)[8]) {
  constexpr int kMaskVal = kMaskedS + (kBiased ? kMagicI : 0);
  if constexpr (kMasked) {
#pragma unroll
    for (int c = 0; c < 128; ++c)
      if (c > lim) s[c] = (uint32_t)kMaskVal;
  }
  // integer max per k-scale class (the bias keeps the order), then 8 dequantised candidates
  int mx[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) mx[k] = kMaskVal;
#pragma unroll
  for (int c = 0; c < 128; ++c) mx[kcls(c)] = max(mx[kcls(c)], (int)s[c]);
  float mrow = -1e30f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    // a class whose columns are all masked must not win the max
    float cand = (kBiased ? (__int_as_float(mx[k]) - kMagicF) : (float)mx[k]) * sc[k];
    if constexpr (kMasked) cand = (mx[k] == kMaskVal) ? -1e30f : cand;
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
#pragma unroll
  for (int k = 0; k < 8; ++k) bias[k] = fmaf(-kMagicF, sc[k], (kFp8 ? kFp8Offset : 0.f) - m_new);
}

// Integer maximum of the 16 columns of k-scale class `k` as a depth-3 tree of 3-input maxima (VIMNMX3): the
// eight classes give 40 independent first-level operations, so a warp that has its scheduler to itself is
// bound by the ALU pipe instead of by an 8-deep dependent chain per class.
template <int kCls>
__device__ __forceinline__ int class_max_tree(const uint32_t (&s)[128]) {
  // columns of class k: block (k >> 2), (c & 7) >> 1 == (k & 3)  ->  c = blk*64 + 8*i + 2*(k&3) + {0,1}
  constexpr int base = (kCls >> 2) * 64 + 2 * (kCls & 3);
  auto e = [&](int n) { return (int)s[base + 8 * (n >> 1) + (n & 1)]; };
  const int m0 = max(max(e(0), e(1)), e(2));
  const int m1 = max(max(e(3), e(4)), e(5));
  const int m2 = max(max(e(6), e(7)), e(8));
  const int m3 = max(max(e(9), e(10)), e(11));
  const int m4 = max(max(e(12), e(13)), e(14));
  const int r0 = max(max(m0, m1), m2);
  const int r1 = max(max(m3, m4), e(15));
  return max(r0, r1);
}

// softmax_max for an unmasked tile with the tree maxima (same results as softmax_max<kFp8, false>).
template <bool kFp8>
__device__ __forceinline__ void softmax_max_tree(const uint32_t (&s)[128], const float (&sc)[8],
                                                 float& m_run, float& alpha, float (&bias)[8]) {
  int mx[8];
  mx[0] = class_max_tree<0>(s); mx[1] = class_max_tree<1>(s); mx[2] = class_max_tree<2>(s); mx[3] = class_max_tree<3>(s);
  mx[4] = class_max_tree<4>(s); mx[5] = class_max_tree<5>(s); mx[6] = class_max_tree<6>(s); mx[7] = class_max_tree<7>(s);
  float cand[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) cand[k] = (float)mx[k] * sc[k];
  const float mrow = fmaxf(fmaxf(fmaxf(cand[0], cand[1]), fmaxf(cand[2], cand[3])),
                           fmaxf(fmaxf(cand[4], cand[5]), fmaxf(cand[6], cand[7])));
  const bool grow = mrow > m_run + (kFp8 ? kFp8Tau : kF16Tau);
  float m_new = m_run;
  alpha = 1.0f;
  if (__any_sync(0xffffffffu, grow)) {
    m_new = fmaxf(m_run, mrow);
    alpha = ex2_approx(m_run - m_new);
  }
  m_run = m_new;
#pragma unroll
  for (int k = 0; k < 8; ++k) bias[k] = fmaf(-kMagicF, sc[k], (kFp8 ? kFp8Offset : 0.f) - m_new);
}

// softmax_max<kFp8, false> in two halves (unmasked tiles): the integer class maxima of columns [64 * kHalf, 64 * kHalf + 64)
// (classes 4 * kHalf .. 4 * kHalf + 3) can be taken as soon as that half of S is in registers; softmax_max_finish then
// gives exactly what softmax_max gives (integer and float maxima do not depend on the order).
template <int kHalf>
__device__ __forceinline__ void softmax_halfmax(const uint32_t (&s)[128], int (&mx)[8]) {
#pragma unroll
  for (int k = 4 * kHalf; k < 4 * kHalf + 4; ++k) mx[k] = kMaskedS;
#pragma unroll
  for (int c = 64 * kHalf; c < 64 * kHalf + 64; ++c) mx[kcls(c)] = max(mx[kcls(c)], (int)s[c]);
}
template <bool kFp8>
__device__ __forceinline__ void softmax_max_finish(const int (&mx)[8], const float (&sc)[8], float& m_run, float& alpha,
                                                   float (&bias)[8]) {
  float mrow = -1e30f;
#pragma unroll
  for (int k = 0; k < 8; ++k) mrow = fmaxf(mrow, (float)mx[k] * sc[k]);
  const bool grow = mrow > m_run + (kFp8 ? kFp8Tau : kF16Tau);
  float m_new = m_run;
  alpha = 1.0f;
  if (__any_sync(0xffffffffu, grow)) {
    m_new = fmaxf(m_run, mrow);
    alpha = ex2_approx(m_run - m_new);
  }
  m_run = m_new;
#pragma unroll
  for (int k = 0; k < 8; ++k) bias[k] = fmaf(-kMagicF, sc[k], (kFp8 ? kFp8Offset : 0.f) - m_new);
}

// 2^y for a pair of logits on the FMA pipe (defined below)
template <int kDeg>
__device__ __forceinline__ uint64_t exp2_poly2(uint64_t y2);
__device__ __forceinline__ uint64_t pack2(float lo, float hi);
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi);

#ifndef SAGE_POLY_NUM
#define SAGE_POLY_NUM 0  // of every SAGE_POLY_DEN column pairs, this many take the polynomial
#endif
#ifndef SAGE_POLY_DEN
#define SAGE_POLY_DEN 4
#endif

__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c);
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b);

#ifndef SAGE_PACKED
#define SAGE_PACKED 1  // dequant FFMA and row-sum FADD as packed f32x2 instructions (half the issue slots); 0: scalar
#endif

// Phase 2: p = 2^(s * sc + bias) in place; returns the row sum of this tile.
template <bool kMasked, int kLo = 0, int kHi = 128, int kDeg = 3, bool kBiased = false>
__device__ __forceinline__ float softmax_exp(uint32_t (&s)[128], const float (&sc)[8],
                                             const float (&bias)[8], int lim) {
  float dsum0 = 0.f, dsum1 = 0.f;
  uint64_t dsum2 = 0;
#pragma unroll
  for (int c = kLo; c < kHi; c += 2) {
    float x0 = __int_as_float((int)s[c] + (kBiased ? 0 : kMagicI));
    float x1 = __int_as_float((int)s[c + 1] + (kBiased ? 0 : kMagicI));
    float y0, y1;
    if (SAGE_PACKED) {
      // columns c and c+1 share their k-scale class
      unpack2(ffma2(pack2(x0, x1), pack2(sc[kcls(c)], sc[kcls(c)]), pack2(bias[kcls(c)], bias[kcls(c)])), y0, y1);
    } else {
      y0 = fmaf(x0, sc[kcls(c)], bias[kcls(c)]);
      y1 = fmaf(x1, sc[kcls(c + 1)], bias[kcls(c + 1)]);
    }
    float p0, p1;
    if (((c >> 1) % SAGE_POLY_DEN) < SAGE_POLY_NUM) {
      unpack2(exp2_poly2<kDeg>(pack2(y0, y1)), p0, p1);
    } else {
      p0 = ex2_approx(y0);
      p1 = ex2_approx(y1);
    }
    if constexpr (kMasked) {
      p0 = (c > lim) ? 0.f : p0;
      p1 = (c + 1 > lim) ? 0.f : p1;
    }
    if (SAGE_PACKED) {
      dsum2 = fadd2(dsum2, pack2(p0, p1));
    } else {
      dsum0 += p0;
      dsum1 += p1;
    }
    s[c] = __float_as_uint(p0);
    s[c + 1] = __float_as_uint(p1);
  }
  if (SAGE_PACKED) unpack2(dsum2, dsum0, dsum1);
  return dsum0 + dsum1;
}

// Phases 2 + 3 fused for one group of 16 columns (kG = 0..7): p = 2^(s * sc + bias), row-sum partials, and
// the group's packed words (4 e4m3 words / 8 f16 words; kG must be a constant after unrolling) - so the conversions (ALU pipe) are issued in the shadow
// of the exponentials (MUFU pipe) instead of after them, and the caller can store the group at once.
template <bool kFp8, bool kMasked>
__device__ __forceinline__ void softmax_exp_pack16(const int kG, uint32_t (&s)[128], const float (&sc)[8],
                                                   const float (&bias)[8], int lim, uint64_t& dsum2,
                                                   uint32_t (&w)[kFp8 ? 4 : 8]) {
#pragma unroll
  for (int i = 0; i < 16; i += 2) {
    const int c = kG * 16 + i;
    const float x0 = __int_as_float((int)s[c] + kMagicI);
    const float x1 = __int_as_float((int)s[c + 1] + kMagicI);
    float y0, y1;
    unpack2(ffma2(pack2(x0, x1), pack2(sc[kcls(c)], sc[kcls(c)]), pack2(bias[kcls(c)], bias[kcls(c)])), y0, y1);
    float p0 = ex2_approx(y0), p1 = ex2_approx(y1);
    if constexpr (kMasked) {
      p0 = (c > lim) ? 0.f : p0;
      p1 = (c + 1 > lim) ? 0.f : p1;
    }
    dsum2 = fadd2(dsum2, pack2(p0, p1));
    s[c] = __float_as_uint(p0);
    s[c + 1] = __float_as_uint(p1);
  }
  if constexpr (kFp8) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int t = kG * 16 + 2 * q;
      w[q] = pack_e4m3x4(__uint_as_float(s[t]), __uint_as_float(s[t + 1]), __uint_as_float(s[t + 8]),
                         __uint_as_float(s[t + 9]));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      w[i] = pack_f16x2(__uint_as_float(s[kG * 16 + 2 * i]), __uint_as_float(s[kG * 16 + 2 * i + 1]));
  }
}

// ------------------------------------------------------------------------------------------
// General attention mask (attn_mask of sageattn_qk_int8_pv_fp16_triton, /root/reference/sageattention/core.py:249-251,
// triton/attn_qk_int8_per_thread.py:39-72).  One path for both kinds, in the float domain:
//   y[c] = s[c] * sc[class] + add[c],   add = 0 / -1e30 for a bool mask (True = attend), or the mask VALUE for a
//   mask of q's dtype - added to the base-2 logits as it is, which is what the reference does (it does not scale
//   the mask by log2 e); columns beyond `lim` (kv_len, as without a mask) count as masked.
// Then the usual lazily-rescaled online softmax on y.  Masked entries contribute exactly 0; a row with nothing
// left keeps m = -1e30, d = 0 (the epilogue writes zeros for it).
// mrow_ptr: the row's mask for THIS kv tile (128 readable elements, 16-byte aligned).  kind: 1 u8 bool, 2 f16, 3 bf16.
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
template <bool kFp8>
__device__ __forceinline__ float softmax_general_mask(uint32_t (&s)[128], const float (&sc)[8], int lim,
                                                      const uint8_t* mrow_ptr, int kind, float& m_run, float& alpha) {
  constexpr float kNeg = -1e30f;
  float mrow = kNeg;
  // pass A: y = dequantised logit + mask, in place (float bits), running maximum
#pragma unroll
  for (int c0 = 0; c0 < 128; c0 += 16) {
    float add[16];
    if (kind == 1) {
      const uint4 w = ldg_nc_v4(mrow_ptr + c0);
      const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 16; ++i) add[i] = ((ww[i >> 2] >> (8 * (i & 3))) & 0xffu) ? 0.f : kNeg;
    } else {
      const uint4 w0 = ldg_nc_v4(mrow_ptr + (size_t)c0 * 2), w1 = ldg_nc_v4(mrow_ptr + (size_t)c0 * 2 + 16);
      const uint32_t ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t h16 = (ww[i >> 1] >> (16 * (i & 1))) & 0xffffu;
        float f;
        if (kind == 2) f = __half2float(__ushort_as_half((unsigned short)h16));
        else f = __uint_as_float(h16 << 16);
        add[i] = fmaxf(f, kNeg);   // -inf (and NaN) -> -1e30: masked
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = c0 + i;
      const float x = __int_as_float((int)s[c] + kMagicI);
      float y = fmaf(x, sc[kcls(c)], -kMagicF * sc[kcls(c)]) + add[i];
      y = (c > lim || add[i] <= kNeg) ? kNeg : y;
      s[c] = __float_as_uint(y);
      mrow = fmaxf(mrow, y);
    }
  }
  const bool grow = mrow > m_run + (kFp8 ? kFp8Tau : kF16Tau);
  float m_new = m_run;
  alpha = 1.0f;
  if (__any_sync(0xffffffffu, grow)) {
    m_new = fmaxf(m_run, mrow);
    alpha = ex2_approx(m_run - m_new);  // 0 on the first unmasked tile (m_run = -1e30)
  }
  m_run = m_new;
  // pass B: p = 2^(y - m [+ offset]); a masked y = -1e30 gives exactly 0 (also while m is still -1e30: the
  // argument is clamped so that -1e30 - (-1e30) does not read as 0)
  const float off = (kFp8 ? kFp8Offset : 0.f) - m_new;
  float d0 = 0.f, d1 = 0.f;
#pragma unroll
  for (int c = 0; c < 128; c += 2) {
    const float y0 = __uint_as_float(s[c]), y1 = __uint_as_float(s[c + 1]);
    const float p0 = (y0 <= kNeg) ? 0.f : ex2_approx(y0 + off);
    const float p1 = (y1 <= kNeg) ? 0.f : ex2_approx(y1 + off);
    d0 += p0;
    d1 += p1;
    s[c] = __float_as_uint(p0);
    s[c + 1] = __float_as_uint(p1);
  }
  return d0 + d1;
}

// Phase 3: pack p to e4m3 (32 words) / f16 (64 words).
template <bool kFp8>
__device__ __forceinline__ void softmax_pack(const uint32_t (&s)[128], uint32_t (&w)[kFp8 ? 32 : 64]) {
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

template <bool kFp8, bool kMasked>
__device__ __forceinline__ void softmax_row(uint32_t (&s)[128], const float (&sc)[8], int lim,
                                            float& m_run, float& d_run, float& alpha,
                                            uint32_t (&w)[kFp8 ? 32 : 64]) {
  float bias[8];
  softmax_max<kFp8, kMasked>(s, sc, lim, m_run, alpha, bias);
  const float dsum = softmax_exp<kMasked>(s, sc, bias, lim);
  d_run = fmaf(d_run, alpha, dsum);
  softmax_pack<kFp8>(s, w);
}

// ------------------------------------------------------------------------------------------
// Half-row forms (64 columns = one 64-row K block = 4 k-scale classes), used by the kernel's
// two-threads-per-row layout.  Same arithmetic as above.
__device__ __forceinline__ constexpr int kcls64(int c) { return (c & 7) >> 1; }

// (masking,) dequantised maximum of the half row; -1e30 when every column is masked.
template <bool kMasked>
__device__ __forceinline__ float half_rowmax(uint32_t (&s)[64], const float (&sc)[4], int lim) {
  if constexpr (kMasked) {
#pragma unroll
    for (int c = 0; c < 64; ++c)
      if (c > lim) s[c] = (uint32_t)kMaskedS;
  }
  int mx[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) mx[k] = kMaskedS;
#pragma unroll
  for (int c = 0; c < 64; ++c) mx[kcls64(c)] = max(mx[kcls64(c)], (int)s[c]);
  float m = -1e30f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float cand = (float)mx[k] * sc[k];
    if constexpr (kMasked) cand = (mx[k] == kMaskedS) ? -1e30f : cand;
    m = fmaxf(m, cand);
  }
  return m;
}

// p = 2^(s * sc + bias) in place; returns the sum.
template <bool kMasked>
__device__ __forceinline__ float half_exp(uint32_t (&s)[64], const float (&sc)[4],
                                          const float (&bias)[4], int lim) {
  float dsum0 = 0.f, dsum1 = 0.f;
#pragma unroll
  for (int c = 0; c < 64; c += 2) {
    float x0 = __int_as_float((int)s[c] + kMagicI);
    float x1 = __int_as_float((int)s[c + 1] + kMagicI);
    float p0 = ex2_approx(fmaf(x0, sc[kcls64(c)], bias[kcls64(c)]));
    float p1 = ex2_approx(fmaf(x1, sc[kcls64(c + 1)], bias[kcls64(c + 1)]));
    if constexpr (kMasked) {
      p0 = (c > lim) ? 0.f : p0;
      p1 = (c + 1 > lim) ? 0.f : p1;
    }
    dsum0 += p0;
    dsum1 += p1;
    s[c] = __float_as_uint(p0);
    s[c + 1] = __float_as_uint(p1);
  }
  return dsum0 + dsum1;
}

// pack to e4m3 (16 words, K index permuted inside 16-groups as in softmax_pack) / f16 (32 words)
template <bool kFp8>
__device__ __forceinline__ void half_pack(const uint32_t (&s)[64], uint32_t (&w)[kFp8 ? 16 : 32]) {
  if constexpr (kFp8) {
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int t = g * 16 + 2 * q;
        w[g * 4 + q] = pack_e4m3x4(__uint_as_float(s[t]), __uint_as_float(s[t + 1]),
                                   __uint_as_float(s[t + 8]), __uint_as_float(s[t + 9]));
      }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      w[i] = pack_f16x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1]));
  }
}

// ------------------------------------------------------------------------------------------
// Streaming form.  The row is processed in four 32-column chunks against the STALE running max
// (exactly the numbers softmax_row produces when no row grows by more than tau), so max tracking,
// exp2, row sum and packing of a chunk are independent instruction streams the scheduler can
// interleave, and chunk c+1 is in flight from TMEM while chunk c is computed.  The caller redoes the
// row when softmax_grew() reports growth beyond tau.  Arithmetic is packed (FFMA2 / FADD2) and a
// compile-time fraction of the exponentials runs as a Cody-Waite polynomial on the FMA pipe.
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};\n" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;\n" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;\n" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;\n" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;\n" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// 2^y for a pair on the FMA pipe: n = round(y), 2^(y-n) by a minimax polynomial on [-0.5, 0.5]
// (relative error 1.7e-3 at degree 2, 7.5e-5 at degree 3), exponent patched in with a shift-add.
template <int kDeg>
__device__ __forceinline__ uint64_t exp2_poly2(uint64_t y2) {
  float y0, y1;
  unpack2(y2, y0, y1);
  y2 = pack2(fmaxf(y0, -125.f), fmaxf(y1, -125.f));
  const uint64_t r2 = fadd2(y2, pack2(kMagicF, kMagicF));
  const uint64_t fl2 = fadd2(r2, pack2(-kMagicF, -kMagicF));
  const uint64_t f2 = ffma2(fl2, pack2(-1.f, -1.f), y2);
  uint64_t p2;
  if constexpr (kDeg == 2) {
    p2 = ffma2(f2, pack2(0.23842119f, 0.23842119f), pack2(0.70343381f, 0.70343381f));
    p2 = ffma2(p2, f2, pack2(1.00044262f, 1.00044262f));
  } else {
    p2 = ffma2(f2, pack2(0.05517084f, 0.05517084f), pack2(0.24260935f, 0.24260935f));
    p2 = ffma2(p2, f2, pack2(0.69326097f, 0.69326097f));
    p2 = ffma2(p2, f2, pack2(0.99992818f, 0.99992818f));
  }
  float r0, r1, p0, p1;
  unpack2(r2, r0, r1);
  unpack2(p2, p0, p1);
  p0 = __int_as_float(__float_as_int(p0) + (__float_as_int(r0) << 23));
  p1 = __int_as_float(__float_as_int(p1) + (__float_as_int(r1) << 23));
  return pack2(p0, p1);
}

struct SoftmaxConsts {
  float sc[8];    // log2e * sm_scale * q_scale * k_scale[class]
  float bias[8];  // offset - m_use - magic * sc[class]
};

template <bool kFp8>
__device__ __forceinline__ void softmax_set_bias(SoftmaxConsts& k, float m_use) {
#pragma unroll
  for (int i = 0; i < 8; ++i) k.bias[i] = fmaf(-kMagicF, k.sc[i], (kFp8 ? kFp8Offset : 0.f) - m_use);
}

// One 32-column chunk (kChunk = 0..3 of the 128-wide tile).  kPolyNum of every kPolyDen column
// pairs take the polynomial.  mx[] = per-class integer maxima, dsum2 = packed partial row sums.
// w: 8 words e4m3 / 16 words f16.  lim: last visible column of the tile (kMasked only).
template <bool kFp8, bool kMasked, int kChunk, int kPolyNum, int kPolyDen, int kDeg>
__device__ __forceinline__ void softmax_chunk(const uint32_t (&s)[32], const SoftmaxConsts& k, int lim,
                                              int (&mx)[8], uint64_t& dsum2, uint32_t (&w)[kFp8 ? 8 : 16]) {
  float p[32];
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    constexpr int kBase = (kChunk >> 1) * 4;
    const int cls = kBase + ((i & 7) >> 1);
    int s0 = (int)s[i], s1 = (int)s[i + 1];
    if constexpr (kMasked) {
      s0 = (kChunk * 32 + i > lim) ? kMaskedS : s0;
      s1 = (kChunk * 32 + i + 1 > lim) ? kMaskedS : s1;
    }
    mx[cls] = max(mx[cls], max(s0, s1));
    const uint64_t x2 = pack2(__int_as_float(s0 + kMagicI), __int_as_float(s1 + kMagicI));
    const uint64_t y2 = ffma2(x2, pack2(k.sc[cls], k.sc[cls]), pack2(k.bias[cls], k.bias[cls]));
    uint64_t p2;
    if (((i >> 1) % kPolyDen) < kPolyNum) {
      p2 = exp2_poly2<kDeg>(y2);
    } else {
      float y0, y1;
      unpack2(y2, y0, y1);
      p2 = pack2(ex2_approx(y0), ex2_approx(y1));
    }
    if constexpr (kMasked) {
      float p0, p1;
      unpack2(p2, p0, p1);
      p0 = (kChunk * 32 + i > lim) ? 0.f : p0;
      p1 = (kChunk * 32 + i + 1 > lim) ? 0.f : p1;
      p2 = pack2(p0, p1);
    }
    dsum2 = fadd2(dsum2, p2);
    unpack2(p2, p[i], p[i + 1]);
  }
  if constexpr (kFp8) {
    // same K-index permutation as softmax_row (word q of a 16-group = tokens 2q, 2q+1, 2q+8, 2q+9)
#pragma unroll
    for (int g = 0; g < 2; ++g)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int t = g * 16 + 2 * q;
        w[g * 4 + q] = pack_e4m3x4(p[t], p[t + 1], p[t + 8], p[t + 9]);
      }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = pack_f16x2(p[2 * i], p[2 * i + 1]);
  }
}

// Row maximum (dequantised) from the per-class integer maxima; -1e30 when every column is masked.
template <bool kMasked>
__device__ __forceinline__ float softmax_rowmax(const int (&mx)[8], const SoftmaxConsts& k) {
  float mrow = -1e30f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float cand = (float)mx[i] * k.sc[i];
    if constexpr (kMasked) cand = (mx[i] == kMaskedS) ? -1e30f : cand;
    mrow = fmaxf(mrow, cand);
  }
  return mrow;
}

}  // namespace sage
