// SageAttention forward for sm_100a (B200), FP16 P.V flavour: INT8 QK^T + online softmax + f16 P.V
//
// Replaces the reference's mma.sync kernel
//   qk_int_sv_f16_attn_kernel  (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671)
// Same from-scratch tcgen05 design as attn_sm100.cu (two 128-row Q tiles per CTA, TMA producer warp,
// single-thread MMA issuer, two softmax warpgroups with one thread per query row, lazy O rescale,
// setmaxnreg 232/40), with ONE difference that was measured to pay for f16 P (4-5 %): P_t is written
// back over the consumed S_t columns in TMEM with tcgen05.st and read by the PV MMA as its TMEM A
// operand, instead of going through a shared-memory tile.  An f16 P tile is 32 KB per Q tile and
// step; through shared memory it costs 64 KB of st.shared plus 64 KB of extra MMA operand reads per
// step pair on a port the tensor pipe already saturates, which outweighs the PV(j) + S(j+1) wait the
// aliasing brings back (issue order PV0(j) S0(j+1) PV1(j) S1(j+1)).  For e4m3 P (half the bytes) the
// balance tips the other way, see attn_sm100.cu and tools/experiments/README.md.
//
// TMEM map (512 columns):  S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D);
// P_t aliases the first 64 columns of S_t.  Numerics contract: as attn_sm100.cu.
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

namespace sage_f16 {
using namespace sage;

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

// 0 = the two softmax warpgroups run freely, 2 = whole softmax steps alternate (see attn_sm100.cu)
#ifndef SAGE_TURN
#define SAGE_TURN 0
#endif

// Development aid: warpgroup 1 starts its first step this many SM cycles late (phase offset
// between the two softmax warpgroups without any per-step barrier).
#ifndef SAGE_DELAY
#define SAGE_DELAY 0
#endif

constexpr int kTileQ = 128;       // rows per Q tile (= TMEM lanes)
constexpr int kTileKV = 128;      // kv rows per step
constexpr int kKStages = 2;
constexpr int kVStages = 2;
constexpr int kSoftmaxThreads = 256;  // 2 warpgroups
constexpr int kThreads = 384;         // + a third warpgroup: TMA warp, MMA warp, two idle warps
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
  static constexpr int kBarOff = kVOff + kVStages * kVBytes;
  static constexpr int kNumBars = 1 + 2 * kKStages + 2 * kVStages + 2 + 2 + 2;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kTotal = kTmemPtrOff + 16;
  static constexpr int kDynamic = kTotal + 1024;  // slack for 1024B alignment
};

// Softmax warpgroup body for Q tile `t` (128 threads, thread = query row).
template <int D, bool kFp8>
__device__ __forceinline__ void softmax_warpgroup(const AttnKernelParams& p, int t, int row,
                                                  int q_tile_start, int n_iter, int b, int h,
                                                  int hkv, uint32_t tmem_base, uint64_t* s_full,
                                                  uint64_t* p_full, uint64_t* o_full, int n_pp) {
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
  if (SAGE_TURN == 2 && t == 1 && n_pp > 0) named_bar_arrive(1, 2 * kTileQ);
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
    if (SAGE_TURN == 2 && j < n_pp) named_bar_sync(1 + t, 2 * kTileQ);
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
    SAGE_STAMP(t, j, 2);
    if (p.debug_s != nullptr && j == 0) {
      const int nq_pad = (p.Nq + 255) / 256 * 256;
      int* dst = p.debug_s + ((size_t)(b * p.Hq + h) * nq_pad + q_row) * 128;
#pragma unroll
      for (int c = 0; c < 128; ++c) dst[c] = (int)s[c];
    }
    const int kv0 = j * kTileKV;
    int lim = p.Nk - 1 - kv0;
    if (p.causal) lim = min(lim, q_row - kv0);
    const bool masked_tile = (j == n_iter - 1) && (p.causal || (p.Nk - kv0) < kTileKV);
    float alpha, bias[8];
    uint32_t w[kFp8 ? 32 : 64];
    if (masked_tile) softmax_max<kFp8, true>(s, sc, lim, m_run, alpha, bias);
    else softmax_max<kFp8, false>(s, sc, 127, m_run, alpha, bias);
    const float dsum = masked_tile ? softmax_exp<true>(s, sc, bias, lim) : softmax_exp<false>(s, sc, bias, 127);
    d_run = fmaf(d_run, alpha, dsum);
    softmax_pack<kFp8>(s, w);
    // rescale O when some row of this warp moved its max (PV(j-1) is complete: S(j) was issued
    // after it and tcgen05.commit tracks all earlier MMAs)
    if (j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {
#pragma unroll
      for (int c0 = 0; c0 < D; c0 += 32) {
        uint32_t o[32];
        tmem_ld32(tO + c0, o);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
        tmem_st32(tO + c0, o);
      }
    }
    if constexpr (kFp8) {
      tmem_st32(tS, w);
    } else {
      uint32_t(&w0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[0]);
      uint32_t(&w1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[32]);
      tmem_st32(tS, w0);
      tmem_st32(tS + 32, w1);
    }
    tc_wait_st();
    tc_fence_before();
    mbar_arrive(&p_full[t]);
    if (SAGE_TURN == 2 && j < n_pp - t) named_bar_arrive(2 - t, 2 * kTileQ);
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
  uint64_t* p_full = s_full + 2;              // 2
  uint64_t* o_full = p_full + 2;              // 2
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
      mbar_init(&p_full[t], kTileQ);
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
                                 s_full, p_full, o_full, min(n_it[0], n_it[1]));
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
          mma_i8_ss(tmem_base + t * 128, ad, bd, idesc_qk, kk > 0);
        }
      };
      auto issue_pv = [&](int t, int j) {
        const uint32_t b0 = sV + (j % kVStages) * L::kVBytes;
        const uint32_t d_tmem = tmem_base + 256 + t * D;
        const uint32_t a_tmem = tmem_base + t * 128;
        if constexpr (kFp8) {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 32; ++kk) {
            const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, 1024, SWZ_128B);
            mma_f8_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, (j > 0) || (kk > 0));
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 16; ++kk) {
            // MN-major B: 16 kv rows x 128 B per MMA = 2 KB step; 64-wide d blocks 16 KB apart
            const uint64_t bd = make_smem_desc(b0 + kk * 2048, kTileKV * 128, 1024, SWZ_128B);
            mma_f16_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, (j > 0) || (kk > 0));
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
        if (n_it[t] > 0) { issue_qk(t, 0); tc_commit(&s_full[t]); }
      tc_commit(&k_empty[0]);

      for (int j = 0; j < n_max; ++j) {
        mbar_wait(&v_full[j % kVStages], (j / kVStages) & 1);
        const bool more = (j + 1 < n_max);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (j < n_it[t]) {
            SAGE_STAMP(2, j, t * 4 + 0);
            mbar_wait(&p_full[t], j & 1);
            tc_fence_after();
            SAGE_STAMP(2, j, t * 4 + 1);
            issue_pv(t, j);
            SAGE_STAMP(2, j, t * 4 + 2);
            if (j == n_it[t] - 1) tc_commit(&o_full[t]);
          }
          if (t == 1) tc_commit(&v_empty[j % kVStages]);
          if (t == 0 && more) {
            mbar_wait(&k_full[(j + 1) % kKStages], ((j + 1) / kKStages) & 1);
            tc_fence_after();
          }
          if (j + 1 < n_it[t]) { issue_qk(t, j + 1); tc_commit(&s_full[t]); }
          SAGE_STAMP(2, j, t * 4 + 3);
          if (t == 1 && more) tc_commit(&k_empty[(j + 1) % kKStages]);
        }
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

}  // namespace sage_f16

// ==========================================================================================
// Host side: TMA descriptors + launch
// ==========================================================================================
namespace sage_f16 {
using namespace sage;

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

}  // namespace sage_f16

extern "C" int sage_attn_i8_f16(const sage_attn_args* a, void* stream) {
  int rc = sage_f16::check_attn_args(a);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return a->D == 128 ? sage_f16::launch_attn<128, false>(a, s) : sage_f16::launch_attn<64, false>(a, s);
}
