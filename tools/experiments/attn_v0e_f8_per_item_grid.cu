// SageAttention forward for sm_100a (B200), FP8 P.V flavour: INT8 QK^T + online softmax + e4m3 P.V
// (sage_attn_i8_f8; the f16 P.V entry point lives in attn_f16_sm100.cu)
//
// Replaces the reference's mma.sync kernels
//   qk_int_sv_f16_attn_kernel  (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671)
//   qk_int_sv_f8_attn_kernel   (/root/reference/csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh:44-707)
// with a from-scratch tcgen05 design (not a port):
//
//   * one CTA owns TWO 128-row Q tiles (256 query rows) of one (batch, head) and walks the KV
//     sequence in 128-row tiles;
//   * warp 8 is the TMA producer (Q once, then a K ring and a V ring, 128B-swizzled smem);
//   * warp 9 is the single-thread MMA issuer:  S_t = Q_t.K^T  as tcgen05.mma kind::i8 (s32 in TMEM),
//     O_t += P_t.V as kind::f8f6f4 (e4m3) or kind::f16, P read from shared memory;
//   * warps 0-3 / 4-7 are two softmax warpgroups, one per Q tile, one thread per query row:
//     tcgen05.ld S (whole row to registers, which frees S_t at once) -> dequant (q/k scales) ->
//     running max / exp2 / row-sum -> P packed to e4m3/f16 and written to a 128B-swizzled smem tile;
//     they also rescale O in TMEM when a row maximum moves by more than tau, and run the epilogue
//     (1/d, v_scale, v_mean, cast, store, LSE).
//   Because P does not alias S, the issuer launches S_t(j+1) as soon as warpgroup t has drained
//   S_t(j): the QK MMA runs under that warpgroup's own softmax and the warpgroups never wait for
//   the tensor pipe in steady state (measured: tools/trace_attn.py).
//
// TMEM map (512 columns):  S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D).
//
// Numerics contract (tolerance-checked against the reference kernels and fp32 SDPA):
//   s' = S * q_scale * k_scale * sm_scale * log2(e);  m = running row max;  p = ex2(s' - m [+8.807]);
//   d = sum of f32 p;  O = sum p.v (f32 in TMEM);  out = O / d [* v_scale] [+ v_mean];
//   lse = log2(d) + m [- 8.807]   (base 2, as the reference kernels return it).
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
#include "softmax_sm100.cuh"
#include "status.h"

namespace sage {

// Development aid (-DSAGE_TRACE, tools/trace_attn.py): SM-clock stamps of CTA (0,0,0)'s pipeline.
#ifdef SAGE_TRACE
constexpr int kTraceSteps = 48, kTraceSlots = 8;
__device__ long long g_trace[3][kTraceSteps][kTraceSlots];
#define SAGE_STAMP(who, j, slot)                                                             \
  do {                                                                                       \
    if (trace_on && (j) < kTraceSteps) g_trace[who][j][slot] = clock64();                    \
  } while (0)
#else
#define SAGE_STAMP(who, j, slot) do { } while (0)
#endif

// How the two softmax warpgroups take turns: 1 = only their exp phases alternate, 2 = whole softmax
// steps alternate (a warpgroup computes alone on its schedulers while the other one's MMAs run).
#ifndef SAGE_TURN
#define SAGE_TURN 1
#endif

// Development aid: warpgroup 1 starts its first step this many SM cycles late (phase offset
// between the two softmax warpgroups without any per-step barrier).
#ifndef SAGE_DELAY
#define SAGE_DELAY 0
#endif

// 1 (measured slower, 14.4 vs 13.4 ms, default 0): the softmax warpgroups pre-set S_t in TMEM to the magic constant 0x4B400000 before every QK MMA
// (which then always accumulates), so the logits come out of tcgen05.ld as the bits of the float
// 1.5*2^23 + S: 8 tcgen05.st per thread and step replace 128 integer adds.
#ifndef SAGE_BIAS_S
#define SAGE_BIAS_S 0
#endif

constexpr int kTileQ = 128;       // rows per Q tile (= TMEM lanes)
constexpr int kTileKV = 128;      // kv rows per step
constexpr int kKStages = 2;
constexpr int kVStages = 2;
constexpr int kSoftmaxThreads = 256;  // 2 warpgroups
constexpr int kThreads = 384;         // + a third warpgroup: TMA warp, MMA warp, two idle warps
// setmaxnreg: the launch gives every warp 168 registers; the third warpgroup hands most of its
// share to the softmax warpgroups, whose exp loop needs the headroom to keep many logits in flight.
constexpr int kSoftmaxRegs = 232;
constexpr int kIssueRegs = 40;
struct AttnKernelParams {
  const float* q_scale;
  const float* k_scale;
  const float* v_scale;  // may be null
  const float* v_mean;   // may be null
  void* o;
  float* lse;  // may be null
  int* debug_s;  // may be null: raw S of kv tile 0, [B,Hq,ceil256(Nq),128] s32
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
  static constexpr int kQBytes = kTileQ * D;                            // int8
  static constexpr int kKBytes = kTileKV * D;                           // int8
  static constexpr int kVBytes = kFp8 ? D * kTileKV : kTileKV * D * 2;  // e4m3 [D][128] | f16 [128][D]
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + 2 * kQBytes;
  static constexpr int kVOff = kKOff + kKStages * kKBytes;
  static constexpr int kPBytes = kFp8 ? kTileQ * kTileKV : kTileQ * kTileKV * 2;  // K-major, SW128
  static constexpr int kPOff = kVOff + kVStages * kVBytes;
  static constexpr int kBarOff = kPOff + 2 * kPBytes;
  static constexpr int kNumBars = 1 + 2 * kKStages + 2 * kVStages + 5 * 2;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kTotal = kTmemPtrOff + 16;
  static constexpr int kDynamic = kTotal + 1024;  // slack for 1024B alignment
};

// Softmax warpgroup body for Q tile `t` (128 threads, thread = query row).
template <int D, bool kFp8>
__device__ __forceinline__ void softmax_warpgroup(const AttnKernelParams& p, int t, int row,
                                                  int q_tile_start, int n_iter, int b, int h,
                                                  int hkv, uint32_t tmem_base, uint8_t* p_smem,
                                                  uint64_t* s_full, uint64_t* s_free,
                                                  uint64_t* p_full, uint64_t* p_empty,
                                                  uint64_t* o_full, int n_pingpong) {
  const int warp_in_wg = (row >> 5);
  const uint32_t lane_off = (uint32_t)(warp_in_wg * 32) << 16;
  const uint32_t tS = tmem_base + lane_off + (uint32_t)(t * 128);
  const uint32_t tO = tmem_base + lane_off + (uint32_t)(256 + t * D);
  const int q_row = q_tile_start + row;

  // per-row q scale (folded with sm_scale * log2 e)
  float qs;
  {
    int idx;
    if (p.gran == 3) idx = (q_row / p.warp_q) * 8 + (q_row & 7);
    else if (p.gran == 2) idx = q_row / p.warp_q;
    else idx = q_row / 128;
    idx = min(idx, p.q_scale_per_bh - 1);
    qs = p.q_scale[(size_t)(b * p.Hq + h) * p.q_scale_per_bh + idx] * p.sm_scale_log2;
  }
  const float* ksp = p.k_scale + (size_t)(b * p.Hkv + hkv) * p.k_scale_per_bh;

  float m_run = -1e30f, d_run = 0.f;
  constexpr bool kBiased = SAGE_BIAS_S != 0;
  auto preset_s = [&]() {
    if constexpr (kBiased) {
      uint32_t mg[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) mg[i] = (uint32_t)kMagicI;
#pragma unroll
      for (int c0 = 0; c0 < 128; c0 += 16) tmem_st16(tS + c0, mg);
      tc_wait_st();
    }
  };
  if constexpr (kBiased) {
    // S_t(0) accumulates onto the constant as well: phase 0 of s_free is "preset done"
    preset_s();
    tc_fence_before();
    mbar_arrive(&s_free[t]);
  }
  // named barrier 1: "warpgroup 0 may start exp" (first pass granted here by warpgroup 1),
  // named barrier 2: "warpgroup 1 may start exp".  Only the first n_pingpong steps, which both
  // warpgroups execute, take part.
  if (SAGE_TURN != 0 && t == 1 && n_pingpong > 0) named_bar_arrive(1, 2 * kTileQ);
#ifdef SAGE_TRACE
  const bool trace_on = (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && row == 0;
#endif

  if (SAGE_DELAY > 0 && t == 1) {
    const long long t_go = clock64() + SAGE_DELAY;
    while (clock64() < t_go) __nanosleep(100);
  }
  for (int j = 0; j < n_iter; ++j) {
    // k scales of this tile -> 8 class scales
    float sc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int idx;
      const int blk64 = j * 2 + (k >> 2);
      if (p.gran == 3) idx = blk64 * 4 + (k & 3);
      else idx = blk64;  // per_warp / per_block: one scale per 64 kv rows
      idx = min(idx, p.k_scale_per_bh - 1);
      sc[k] = __ldg(ksp + idx) * qs;
    }
    SAGE_STAMP(t, j, 0);
    mbar_wait(&s_full[t], j & 1);
    if (SAGE_TURN == 2 && j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
    tc_fence_after();
    SAGE_STAMP(t, j, 1);
    uint32_t s[128];
    {
      uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
      uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
      uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
      uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
      tmem_ld32(tS + 0, s0);
      tmem_ld32(tS + 32, s1);
      tmem_ld32(tS + 64, s2);
      tmem_ld32(tS + 96, s3);
      tc_wait_ld();
    }
    // S_t is in registers: the issuer may overwrite it with S_t(j+1)
    preset_s();
    tc_fence_before();
    mbar_arrive(&s_free[t]);
    SAGE_STAMP(t, j, 2);
    if (p.debug_s != nullptr && j == 0) {
      const int nq_pad = (p.Nq + 255) / 256 * 256;
      int* dst = p.debug_s + ((size_t)(b * p.Hq + h) * nq_pad + q_row) * 128;
#pragma unroll
      for (int c = 0; c < 128; ++c) dst[c] = (int)s[c] - (kBiased ? kMagicI : 0);
    }
    const int kv0 = j * kTileKV;
    int lim = p.Nk - 1 - kv0;
    if (p.causal) lim = min(lim, q_row - kv0);
    const bool masked_tile = (j == n_iter - 1) && (p.causal || (p.Nk - kv0) < kTileKV);
    float alpha, bias[8], dsum;
    uint32_t w[kFp8 ? 32 : 64];
    if (masked_tile) softmax_max<kFp8, true, kBiased>(s, sc, lim, m_run, alpha, bias);
    else softmax_max<kFp8, false, kBiased>(s, sc, 127, m_run, alpha, bias);
    SAGE_STAMP(t, j, 5);
    auto rescale_o = [&]() {
      // PV(j-1) complete: P tile reusable and O_t stable
      if (j > 0) {
        mbar_wait(&p_empty[t], (j - 1) & 1);
        // rescale O when some row of this warp moved its max
        if (__any_sync(0xffffffffu, alpha != 1.0f)) {
          tc_fence_after();
#pragma unroll
          for (int c0 = 0; c0 < D; c0 += 32) {
            uint32_t o[32];
            tmem_ld32(tO + c0, o);
            tc_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(tO + c0, o);
          }
          tc_wait_st();
          tc_fence_before();
        }
      }
    };
    auto store_p = [&]() {
      // P row -> smem, K-major 128B-swizzled (16B chunk c of row r sits at chunk c ^ (r & 7))
      const uint32_t prow = smem_u32(p_smem) + (uint32_t)row * 128u;
      const uint32_t x = (uint32_t)(row & 7);
#pragma unroll
      for (int c = 0; c < (kFp8 ? 8 : 16); ++c) {
        const uint32_t addr = prow + (uint32_t)(c >> 3) * (kTileQ * 128) + (((uint32_t)(c & 7) ^ x) << 4);
        st_shared_v4(addr, w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
      }
    };
    if (SAGE_TURN == 3) {
      // The MUFU-bound stretch of the two warpgroups alternates, and it is ONE straight-line block:
      // exponentials, row sum, packing and the smem stores, so that the conversions and stores of
      // finished columns fill the issue slots the MUFU cadence leaves (the O-rescale check, which
      // only needs alpha, moves in front of it).
      rescale_o();
      if (j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
      SAGE_STAMP(t, j, 6);
      if (masked_tile) dsum = softmax_exp<true, 0, 128, 3, kBiased>(s, sc, bias, lim);
      else dsum = softmax_exp<false, 0, 128, 3, kBiased>(s, sc, bias, 127);
      d_run = fmaf(d_run, alpha, dsum);
      softmax_pack<kFp8>(s, w);
      store_p();
      if (j < n_pingpong - t) named_bar_arrive(2 - t, 2 * kTileQ);
      SAGE_STAMP(t, j, 7);
      SAGE_STAMP(t, j, 3);
    } else {
      // The exp phases of the two warpgroups alternate (they are the only MUFU-bound stretch): each
      // warpgroup's load / max / pack / store phases then run under the other one's exponentials.
      if (SAGE_TURN == 1 && j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
      SAGE_STAMP(t, j, 6);
      if (masked_tile) dsum = softmax_exp<true, 0, 128, 3, kBiased>(s, sc, bias, lim);
      else dsum = softmax_exp<false, 0, 128, 3, kBiased>(s, sc, bias, 127);
      if (SAGE_TURN == 1 && j < n_pingpong - t) named_bar_arrive(2 - t, 2 * kTileQ);
      SAGE_STAMP(t, j, 7);
      d_run = fmaf(d_run, alpha, dsum);
      softmax_pack<kFp8>(s, w);
      SAGE_STAMP(t, j, 3);
      rescale_o();
      store_p();
    }
    fence_proxy_async_smem();
    mbar_arrive(&p_full[t]);
    if (SAGE_TURN == 2 && j < n_pingpong - t) named_bar_arrive(2 - t, 2 * kTileQ);
    SAGE_STAMP(t, j, 4);
  }

  // ---- epilogue
  mbar_wait(&o_full[t], 0);
  tc_fence_after();
  const float inv_d = rcp_approx(d_run);
  const bool store = q_row < p.Nq;
  const float* vsp = p.v_scale ? p.v_scale + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  const float* vmp = p.v_mean ? p.v_mean + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  uint16_t* optr = reinterpret_cast<uint16_t*>(p.o) + (size_t)b * p.o_stride_b +
                   (size_t)h * p.o_stride_h + (size_t)q_row * p.o_stride_n;
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 32) {
    uint32_t o[32];
    tmem_ld32(tO + c0, o);
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
        lg2_approx(d_run) + m_run - (kFp8 ? kFp8Offset : 0.f);
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
  uint64_t* s_full = v_empty + kVStages;      // 2
  uint64_t* s_free = s_full + 2;              // 2
  uint64_t* p_full = s_free + 2;              // 2
  uint64_t* p_empty = p_full + 2;             // 2
  uint64_t* o_full = p_empty + 2;             // 2
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);

  const int warp = threadIdx.x >> 5;
  // heavy (late) q tiles first: helps the causal tail
  const int q_pair = (int)gridDim.x - 1 - (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hkv = h / (p.Hq / p.Hkv);
  const int q_start = q_pair * 2 * kTileQ;

  // kv tiles per q tile (causal: stop at the diagonal tile; top-left aligned as the reference,
  // /root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:218-222)
  int n_it[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int qs = q_start + t * kTileQ;
    int kv_end = p.Nk;
    if (p.causal) kv_end = min(kv_end, qs + kTileQ);
    n_it[t] = (qs < p.Nq) ? (kv_end + kTileKV - 1) / kTileKV : 0;
  }
  const int n_max = max(n_it[0], n_it[1]);

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < kKStages; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < kVStages; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1);
      mbar_init(&s_free[t], kTileQ);
      mbar_init(&p_full[t], kTileQ);
      mbar_init(&p_empty[t], 1);
      mbar_init(&o_full[t], 1);
    }
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
                                 smem + L::kPOff + t * L::kPBytes, s_full, s_free, p_full, p_empty,
                                 o_full, min(n_it[0], n_it[1]));
  } else if (warp == 8) {
    // ======================= TMA producer =======================
    reg_dealloc<kIssueRegs>();
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, 2 * L::kQBytes);
      tma_load_4d(smem + L::kQOff, &tmQ, q_full, 0, q_start, h, b);
      tma_load_4d(smem + L::kQOff + L::kQBytes, &tmQ, q_full, 0, q_start + kTileQ, h, b);
      for (int j = 0; j < n_max; ++j) {
        const int ks = j % kKStages, vs = j % kVStages;
        mbar_wait(&k_empty[ks], ((j / kKStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&k_full[ks], L::kKBytes);
        tma_load_4d(smem + L::kKOff + ks * L::kKBytes, &tmK, &k_full[ks], 0, j * kTileKV, hkv, b);
        mbar_wait(&v_empty[vs], ((j / kVStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&v_full[vs], L::kVBytes);
        uint8_t* vdst = smem + L::kVOff + vs * L::kVBytes;
        if constexpr (kFp8) {
          // V^T e4m3 [D rows][128 kv] (kv contiguous)
          tma_load_4d(vdst, &tmV, &v_full[vs], j * kTileKV, 0, hkv, b);
        } else {
          // V f16 [128 kv][D]: one 64-wide (128 B) column block per load
#pragma unroll
          for (int c = 0; c < D / 64; ++c)
            tma_load_4d(vdst + c * (kTileKV * 128), &tmV, &v_full[vs], c * 64, j * kTileKV, hkv, b);
        }
      }
    }
  } else if (warp == 9) {
    // ======================= MMA issuer =======================
    reg_dealloc<kIssueRegs>();
    if (elect_one()) {
      constexpr uint32_t idesc_qk = make_idesc(CFMT_S32, FMT_S8, FMT_S8, 0, 0, kTileQ, kTileKV);
      constexpr uint32_t idesc_pv =
          kFp8 ? make_idesc(CFMT_F32, FMT_E4M3, FMT_E4M3, 0, 0, kTileQ, D)
               : make_idesc(CFMT_F32, FMT_F16, FMT_F16, 0, /*B MN-major*/ 1, kTileQ, D);
      // Q/K rows are D bytes: one 128B (D=128) or 64B (D=64) swizzle span; 8-row groups are dense
      constexpr uint64_t qk_swz = (D == 128) ? SWZ_128B : SWZ_64B;
      constexpr uint32_t qk_sbo = 8 * D;
      const uint32_t sQ = smem_u32(smem + L::kQOff);
      const uint32_t sK = smem_u32(smem + L::kKOff);
      const uint32_t sV = smem_u32(smem + L::kVOff);

      auto issue_qk = [&](int t, int j) {
        const uint32_t a0 = sQ + t * L::kQBytes;
        const uint32_t b0 = sK + (j % kKStages) * L::kKBytes;
#pragma unroll
        for (int kk = 0; kk < D / 32; ++kk) {
          const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, qk_sbo, qk_swz);
          const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, qk_sbo, qk_swz);
          mma_i8_ss(tmem_base + t * 128, ad, bd, idesc_qk, (SAGE_BIAS_S != 0) || kk > 0);
        }
      };
      const uint32_t sP = smem_u32(smem + L::kPOff);
      auto issue_pv = [&](int t, int j) {
        const uint32_t a0 = sP + t * L::kPBytes;
        const uint32_t b0 = sV + (j % kVStages) * L::kVBytes;
        const uint32_t d_tmem = tmem_base + 256 + t * D;
        if constexpr (kFp8) {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 32; ++kk) {
            const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, 1024, SWZ_128B);
            const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, 1024, SWZ_128B);
            mma_f8_ss(d_tmem, ad, bd, idesc_pv, (j > 0) || (kk > 0));
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 16; ++kk) {
            // A: two 64-wide K blocks of [128 rows][128 B], 16 KB apart, 32 B per MMA inside a block
            const uint64_t ad = make_smem_desc(a0 + (kk >> 2) * (kTileQ * 128) + (kk & 3) * 32, 16, 1024, SWZ_128B);
            // MN-major B: 16 kv rows x 128 B per MMA = 2 KB step; 64-wide d blocks 16 KB apart
            const uint64_t bd = make_smem_desc(b0 + kk * 2048, kTileKV * 128, 1024, SWZ_128B);
            mma_f16_ss(d_tmem, ad, bd, idesc_pv, (j > 0) || (kk > 0));
          }
        }
      };

#ifdef SAGE_TRACE
      const bool trace_on = (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
#endif
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      for (int t = 0; t < 2; ++t)
        if (n_it[t] > 0) {
          if (SAGE_BIAS_S != 0) {  // S_t has been pre-set by its warpgroup
            mbar_wait(&s_free[t], 0);
            tc_fence_after();
          }
          issue_qk(t, 0);
          tc_commit(&s_full[t]);
        }
      tc_commit(&k_empty[0]);

      for (int j = 0; j < n_max; ++j) {
        // S(j+1) as soon as the warpgroup has drained S(j)
        if (j + 1 < n_max) {
          mbar_wait(&k_full[(j + 1) % kKStages], ((j + 1) / kKStages) & 1);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (j + 1 < n_it[t]) {
              SAGE_STAMP(2, j, t * 4 + 0);
              mbar_wait(&s_free[t], (j + (SAGE_BIAS_S != 0 ? 1 : 0)) & 1);
              tc_fence_after();
              SAGE_STAMP(2, j, t * 4 + 1);
              issue_qk(t, j + 1);
              tc_commit(&s_full[t]);
            }
          }
          tc_commit(&k_empty[(j + 1) % kKStages]);
        }
        mbar_wait(&v_full[j % kVStages], (j / kVStages) & 1);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (j < n_it[t]) {
            SAGE_STAMP(2, j, t * 4 + 2);
            mbar_wait(&p_full[t], j & 1);
            tc_fence_after();
            SAGE_STAMP(2, j, t * 4 + 3);
            issue_pv(t, j);
            tc_commit(&p_empty[t]);
            if (j == n_it[t] - 1) tc_commit(&o_full[t]);
          }
        }
        tc_commit(&v_empty[j % kVStages]);
      }
    }
  } else {
    reg_dealloc<kIssueRegs>();
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
                 a->k8.stride_n, a->k8.stride_h, a->k8.stride_b, D, kTileKV, qk_swz, "k_int8");
  if (rc) return rc;
  if constexpr (kFp8) {
    // v: e4m3 [B,Hkv,D,Npad] rows; tensor4 strides are reused as (b, h, d-row)
    SAGE_REQUIRE(a->npad_v >= a->Nk && a->npad_v % 64 == 0, "v_fp8 padded length %d invalid for kv_len %d", a->npad_v, a->Nk);
    rc = make_map4(&tmV, CU_TENSOR_MAP_DATA_TYPE_UINT8, a->v.ptr, a->npad_v, D, a->Hkv, a->B,
                   a->v.stride_n, a->v.stride_h, a->v.stride_b, kTileKV, D,
                   CU_TENSOR_MAP_SWIZZLE_128B, "v_fp8");
  } else {
    rc = make_map4(&tmV, CU_TENSOR_MAP_DATA_TYPE_UINT16, a->v.ptr, D, a->Nk, a->Hkv, a->B,
                   a->v.stride_n * 2, a->v.stride_h * 2, a->v.stride_b * 2, 64, kTileKV,
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

#ifdef SAGE_TRACE
extern "C" int sage_debug_trace(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, sage::g_trace, sizeof(sage::g_trace));
}
#endif

extern "C" int sage_attn_i8_f8(const sage_attn_args* a, void* stream) {
  int rc = sage::check_attn_args(a);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? sage::launch_attn<128, true>(a, s) : sage::launch_attn<64, true>(a, s);
}
