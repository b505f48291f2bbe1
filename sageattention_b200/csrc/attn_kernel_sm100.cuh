// SageAttention forward for sm_100a (B200): INT8 QK^T + online softmax + P.V in f16 or e4m3, ONE kernel template
//   sage_attn_fwd_kernel<D, kFp8, kMask, kVarlen>     D = 64 | 128, P.V in e4m3 (kFp8) or f16
// instantiated by attn_f16_sm100.cu (sage_attn_i8_f16) and attn_f8_sm100.cu (sage_attn_i8_f8).
//
// Replaces the reference's mma.sync / wgmma kernels
//   qk_int_sv_f16_attn_kernel  (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671)
//   qk_int_sv_f8_attn_kernel   (/root/reference/csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh:44-..., ..._sm90.cu)
// with a from-scratch tcgen05 design: persistent grid (one CTA per SM) with a dynamic work scheduler, two 128-row Q
// tiles per work item, a TMA producer warp, a single-thread MMA issuer, two softmax warpgroups with one thread per
// query row, lazy O rescale, setmaxnreg 232/40, TMA-store epilogue.
//
// TMEM map (512 columns):  S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D).
// P_t is written back over the consumed S_t columns with tcgen05.st (f16: columns [0,64), e4m3: [0,32)) and read by the
// P.V MMA as its TMEM A operand.  The P -> S dependency this aliasing creates (S_t(j+1) cannot overtake the P.V that
// still reads P_t(j); the tensor pipe executes in order) is taken off the critical path in two ways (round 2, measured
// +8..14 % on every configuration; see SAGE_PSPLIT / SAGE_SSPLIT below and tools/experiments/README.md):
//   * P goes out in two parts, so most of P.V(j) runs under the tail of the exp window;
//   * S(j+1) is issued as two N = 64 halves, the upper one ahead of P.V(j), and the warpgroup works on it while the
//     lower one is still being computed.
// Until then the e4m3 and head_dim-64 flavours used a second kernel (P through shared memory / spare TMEM columns, exp
// windows of the two warpgroups alternating; attn_sm100.cu in the history) - this one now beats it everywhere.
//
// Numerics contract: logits dequantised as s * (q_scale * k_scale[class] * sm_scale * log2 e); online softmax in base 2
// with a lazily rescaled accumulator (tau = 8 for f16 P, 2 for e4m3 P, whose exponent carries the reference's +8.807
// offset, attn_utils.cuh:30); P rounded to f16 / e4m3 (rn, satfinite), P.V accumulated in fp32 in TMEM over the WHOLE
// kv sequence (the reference's sm89 kernel accumulates in fp32 across 32 MMAs and then in registers; sm90 in fp32).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <type_traits>

#include "attn_common_sm100.cuh"

namespace sage_attn {
using namespace sage;

// The two softmax warpgroups run freely here (alternating their exp phases was measured slower with P
// in TMEM).
#ifndef SAGE_TURN
#define SAGE_TURN 0
#endif
#ifndef SAGE_MAXTREE
#define SAGE_MAXTREE 0
#endif
// 1: the k scales of a kv step as two 16-byte loads (per_thread) / two scalars (per_warp) instead of eight indexed loads
#ifndef SAGE_KSVEC
#define SAGE_KSVEC 1
#endif
// f16 / e4m3 conversions issued group by group (16 columns) inside the exp phase, under the MUFU pipe.
// Measured at f16 hd128 N=8192: 1183 (median; 1252 best, noisy) -> 1263 TOPS (stable).  It is also what lets P go out
// in parts (SAGE_PSPLIT).
#ifndef SAGE_FUSE
#define SAGE_FUSE 1
#endif
// 1: S_t(j+1) is issued in two N = 64 halves, each with its own barrier.  An f16 P row occupies only columns [0, 64) of
// the consumed S_t, so the UPPER half of the next S (keys 64..127 -> columns [64, 128)) goes out as soon as the warpgroup
// has pulled the upper half of S_t(j) into registers (s_free) - ahead of P.V(j) on the tensor pipe - and at the start of
// step j+1 the warpgroup finds it ready: it loads it and takes its class maxima while P.V(j) (late part) and the lower
// half of Q.K^T(j+1) are still running.  (Round 2's first form of this waited for both warpgroups' s_free before any
// P.V and had one barrier for the whole S: -1 %.)
#ifndef SAGE_SSPLIT
#define SAGE_SSPLIT 1
#endif

// n > 0: the P row goes out in two parts.  The first n groups of 16 columns are stored (and announced on p_part) while
// the exponentials of the remaining groups are still running, so the first n K-steps of P.V(j) execute under the exp
// window instead of after it; the rest follows on p_full as before.  The lazy O rescale then has to happen BEFORE the
// window (its decision is known after the max phase).  SAGE_PSPLIT_LAG: groups computed between the early tcgen05.st and
// the wait::st + arrive that publishes it.  Measured (burst TOPS, B4 H32, profiles/r2k_*): f16 hd128 N = 2K / 4K / 8K /
// 8K causal 1140 / 1232 / 1272 / 1200 -> 1281 / 1392 / 1446 / 1325 with n = 6 and SAGE_SSPLIT; e4m3 hd128 (against the
// former shared-memory-P kernel) 1169 / - / 1305 / 1217 -> 1257 / - / 1459 / 1335 with n = 4; hd64 e4m3 679 -> 738,
// f16 678 -> 700.  Outputs are bit-identical to the unsplit kernel.
#ifndef SAGE_PSPLIT
#define SAGE_PSPLIT 6
#endif
#ifndef SAGE_PSPLIT_LAG
#define SAGE_PSPLIT_LAG 0
#endif

// 1: the step works through S in two halves, UPPER half first (softmax_steps_spec below): its exponentials are computed
// against the running maximum the row already has - which is what the lazily rescaled softmax would use anyway unless
// the row's maximum grows by more than tau in this tile - while the tensor pipe is still busy with P.V(j-1) (late part)
// and the lower half of Q.K^T(j).  When the lower half is in, the step decides as before; in the rare step where a row of
// the warp does grow, the upper half is pulled from TMEM again (it is still there: s_free is only signalled after the
// decision) and redone.  P then goes out as [upper half | lower half], and P.V(j) is issued in that order of K-steps.
#ifndef SAGE_SPEC
#define SAGE_SPEC 0
#endif
// 2: one MMA issuer warp per Q tile (warps 9 and 10) instead of one for both.  A single in-order issuer serves the
// barriers of warpgroup 0 and warpgroup 1 alternately, so a P part of one warpgroup that is ready waits while the issuer
// is blocked on the other warpgroup's next barrier; with two issuers each warpgroup's MMAs go out as soon as its own
// barriers allow.  The K / V / Q ring slots are then released by two commits (one per issuer).
#ifndef SAGE_ISSUERS
#define SAGE_ISSUERS 1
#endif
#ifndef SAGE_PSPLIT_SMALL
#define SAGE_PSPLIT_SMALL 4   // the e4m3 flavour and head_dim 64 (shorter P.V): measured best at 4
#endif
template <int D, bool kFp8>
__host__ __device__ constexpr int psplit_groups() { return (kFp8 || D == 64) ? SAGE_PSPLIT_SMALL : SAGE_PSPLIT; }

// words [kBegin, kEnd) of a packed P row -> TMEM columns tP + kBegin.. (multiples of 8 words)
template <int kBegin, int kEnd, int kN>
__device__ __forceinline__ void tmem_st_words(uint32_t tP, uint32_t (&w)[kN]) {
  static_assert(kBegin % 8 == 0 && kEnd % 8 == 0 && kEnd <= kN, "8-word units");
  if constexpr (kEnd - kBegin >= 32) {
    tmem_st32(tP + kBegin, *reinterpret_cast<uint32_t(*)[32]>(&w[kBegin]));
    tmem_st_words<kBegin + 32, kEnd>(tP, w);
  } else if constexpr (kEnd - kBegin >= 16) {
    tmem_st16(tP + kBegin, *reinterpret_cast<uint32_t(*)[16]>(&w[kBegin]));
    tmem_st_words<kBegin + 16, kEnd>(tP, w);
  } else if constexpr (kEnd - kBegin >= 8) {
    tmem_st8(tP + kBegin, *reinterpret_cast<uint32_t(*)[8]>(&w[kBegin]));
    tmem_st_words<kBegin + 8, kEnd>(tP, w);
  }
}

template <int D, bool kFp8>
struct SmemLayout {
  static constexpr int kQBytes = kTileQ * D;                            // int8
  static constexpr int kKBytes = kTileKV * D;                           // int8
  static constexpr int kVBytes = kFp8 ? D * kTileKV : kTileKV * D * 2;  // e4m3 [D][128] | f16 [128][D]
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + kQStages * 2 * kQBytes;
  static constexpr int kVOff = kKOff + kKStages * kKBytes;
  static constexpr int kPBytes = 0;  // P lives in TMEM
  static constexpr int kPOff = kVOff + kVStages * kVBytes;
  static constexpr int kOBytes = kTileQ * D * 2;  // output staging tile per warpgroup: D/64 blocks of [128][128 B], SW128
  static constexpr int kOOff = kPOff + 2 * kPBytes;
  static constexpr int kBarOff = kOOff + 2 * kOBytes;
  static constexpr int kNumBars = 2 * kQStages + 2 * kKStages + 2 * kVStages + 6 * 2 + 2 * 3 /*work ring*/;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kRingOff = kTmemPtrOff + 16;  // work ring: kSchedSlots x 16 B
  static constexpr int kVsOff = kRingOff + 4 * 16;    // v_scale [D] f32 of every ring slot (bulk-copied by the producer)
  static constexpr int kTotal = kVsOff + 3 * 128 * 4;
  static constexpr int kDynamic = kTotal + 1024;  // slack for 1024B alignment
};

// The kv steps of one work item for Q tile `t` (128 threads, thread = query row).
template <int D, bool kFp8, bool kMask, bool kVarlen = false>
__device__ __forceinline__ void softmax_steps(const AttnKernelParams& p, int t, int row,
                                              const ItemCtx& c, const WorkItem* ring, uint32_t it,
                                              uint32_t tmem_base, uint64_t* s_full, uint64_t* s_free, uint64_t* p_full,
                                              uint64_t* p_part, uint64_t* s_hi, WgState& st, float& m_out, float& d_out) {
  const int warp_in_wg = (row >> 5);
  const uint32_t lane_off = (uint32_t)(warp_in_wg * 32) << 16;
  const uint32_t tS = tmem_base + lane_off + (uint32_t)(t * 128);
  const uint32_t tO = tmem_base + lane_off + (uint32_t)(256 + t * D);
  const int q_row = c.q_row;
  const float qs = c.qs * p.sm_scale_log2;  // folded with sm_scale * log2 e
  const float* ksp = c.ksp;
  const int n_iter = c.n_iter, n_pingpong = c.n_pingpong;
  const bool next_pingpong = c.next_pingpong;

  float m_run = -1e30f, d_run = 0.f;
  // named barrier 1: "warpgroup 0 may start exp", named barrier 2: "warpgroup 1 may start exp".
  // Only the first n_pingpong steps of an item, which both warpgroups execute, take part.  Warpgroup 1
  // grants warpgroup 0's first turn of an item either here or, when it can, right after its own last
  // exp phase of the previous item (so that warpgroup 0 does not wait for warpgroup 1's epilogue).
  constexpr int kTurn = kMask ? 0 : SAGE_TURN;
  if (kTurn != 0 && t == 1 && n_pingpong > 0) {
    if (!st.granted) named_bar_arrive(1, 2 * kTileQ);
    st.granted = false;
  }
#ifdef SAGE_TRACE
  const bool trace_on = blockIdx.x == 0 && row == 0;
#endif

  for (int j = 0; j < n_iter; ++j) {
    const uint32_t ph = st.steps & 1u;
    // k scales of this tile -> 8 class scales
    float sc[8];
    if constexpr (!kVarlen) {
      if (SAGE_KSVEC && p.gran == 3) {
        // per_thread: the 8 scales of kv tile j are 8 consecutive floats (2 blocks of 64 rows x 4 groups): two 16-byte
        // loads instead of eight indexed ones.  A second block beyond kv_len (ragged last tile) has no scales: all its
        // columns are masked, any finite factor will do - the first block's are reused
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(ksp + j * 8));
        const float4 s1 = (j * 8 + 4 < p.k_scale_per_bh) ? __ldg(reinterpret_cast<const float4*>(ksp + j * 8 + 4)) : s0;
        sc[0] = s0.x * qs; sc[1] = s0.y * qs; sc[2] = s0.z * qs; sc[3] = s0.w * qs;
        sc[4] = s1.x * qs; sc[5] = s1.y * qs; sc[6] = s1.z * qs; sc[7] = s1.w * qs;
      } else if (SAGE_KSVEC && p.gran != 3) {
        // per_warp / per_block: one scale per 64 kv rows
        const float a0 = __ldg(ksp + min(j * 2, p.k_scale_per_bh - 1)) * qs;
        const float a1 = __ldg(ksp + min(j * 2 + 1, p.k_scale_per_bh - 1)) * qs;
        sc[0] = sc[1] = sc[2] = sc[3] = a0;
        sc[4] = sc[5] = sc[6] = sc[7] = a1;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          int idx;
          const int blk64 = j * 2 + (k >> 2);
          if (p.gran == 3) idx = blk64 * 4 + (k & 3);
          else idx = blk64;  // per_warp / per_block: one scale per 64 kv rows
          idx = min(idx, p.k_scale_per_bh - 1);
          sc[k] = __ldg(ksp + idx) * qs;
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        int idx = j * 2 + (k >> 2);             // per-block scales of the sequence, [blocks, H]
        idx = max(min(idx, c.ks_count - 1), 0);   // (0 blocks: the key-less sequence's one masked step)
        sc[k] = __ldg(ksp + (size_t)idx * c.ks_stride) * qs;
      }
    }
    const int kv0 = j * kTileKV;
    const int k_len = kVarlen ? c.k_len : p.Nk;
    int lim = k_len - 1 - kv0;
    if (p.causal) lim = min(lim, q_row - kv0);
    const bool masked_tile = (j == n_iter - 1) && (p.causal || (k_len - kv0) < kTileKV);
    SAGE_STAMP(t, st.steps, 0);
    uint32_t s[128];
    int mx[8];
    {
      uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
      uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
      uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
      uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
      if (SAGE_SSPLIT) {
        mbar_wait(&s_hi[t], ph);
        tc_fence_after();
        tmem_ld32(tS + 64, s2);
        tmem_ld32(tS + 96, s3);
        tc_wait_ld();
        // the upper 64 columns of S_t are in registers and P will not use them: free for the next S
        tc_fence_before();
        mbar_arrive(&s_free[t]);
        if (!kMask && !masked_tile) softmax_halfmax<1>(s, mx);
        mbar_wait(&s_full[t], ph);
        tc_fence_after();
        SAGE_STAMP(t, st.steps, 1);
        tmem_ld32(tS + 0, s0);
        tmem_ld32(tS + 32, s1);
        tc_wait_ld();
      } else {
        mbar_wait(&s_full[t], ph);
        tc_fence_after();
        SAGE_STAMP(t, st.steps, 1);
        tmem_ld32(tS + 0, s0);
        tmem_ld32(tS + 32, s1);
        tmem_ld32(tS + 64, s2);
        tmem_ld32(tS + 96, s3);
        tc_wait_ld();
      }
    }
    SAGE_STAMP(t, st.steps, 2);
    if (p.debug_s != nullptr && j == 0) {
      const int nq_pad = (p.Nq + 255) / 256 * 256;
      const int4 wv = ld_shared_v4(smem_u32(&ring[it % kSchedSlots]));
      WorkItem wi; wi.h = wv.y; wi.b = wv.z;
      int* dst = p.debug_s + ((size_t)(wi.b * p.Hq + wi.h) * nq_pad + q_row) * 128;
#pragma unroll
      for (int c = 0; c < 128; ++c) dst[c] = (int)s[c];
    }
    float alpha, bias[8], dsum;
    uint32_t w[kFp8 ? 32 : 64];
    if constexpr (kMask) {
      // attn_mask: max, exponentials and row sum of y = logit + mask in one call (softmax_sm100.cuh)
      dsum = softmax_general_mask<kFp8>(s, sc, lim, c.mask_row + (size_t)kv0 * (p.mask_kind == 1 ? 1 : 2), p.mask_kind,
                                        m_run, alpha);
    } else if (masked_tile) softmax_max<kFp8, true, false>(s, sc, lim, m_run, alpha, bias);
    else if (SAGE_SSPLIT) {
      softmax_halfmax<0>(s, mx);
      softmax_max_finish<kFp8>(mx, sc, m_run, alpha, bias);
    } else if (SAGE_MAXTREE) softmax_max_tree<kFp8>(s, sc, m_run, alpha, bias);
    else softmax_max<kFp8, false, false>(s, sc, 127, m_run, alpha, bias);
    // rescale O when some row of this warp moved its max (PV(j-1) is complete: S(j) was issued after
    // it and tcgen05.commit tracks all earlier MMAs)
    auto rescale_o = [&]() {
      if (j > 0) {
        if (__any_sync(0xffffffffu, alpha != 1.0f)) {
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
      }
    };
    constexpr int kPs = (kMask || SAGE_FUSE != 1) ? 0 : psplit_groups<D, kFp8>();   // groups in the early part of P
    constexpr int kWg = kFp8 ? 4 : 8;                                  // packed words per 16-column group
    static_assert(kPs >= 0 && kPs < 8 && (!kFp8 || kPs % 2 == 0), "SAGE_PSPLIT");
    if constexpr (kPs > 0) rescale_o();   // before any part of P.V(j) may start
    SAGE_STAMP(t, st.steps, 5);
    // The exp phases of the two warpgroups alternate (they are the only MUFU-bound stretch): each
    // warpgroup's load / max / pack / store phases then run under the other one's exponentials.
    if (kTurn == 1 && j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
    SAGE_STAMP(t, st.steps, 6);
    constexpr int kFuse = kMask ? 0 : SAGE_FUSE;
    if constexpr (kMask) {
      // (exponentials done above)
    } else {
#if SAGE_FUSE
    // conversions issued group by group (16 columns) under the MUFU pipe; SAGE_FUSE == 2 also stores every group
    // over the consumed S columns as soon as it is packed
    auto window = [&](auto masked_c) {
      constexpr bool kM = decltype(masked_c)::value;
      uint64_t dsum2 = 0;
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        constexpr int kW = kFp8 ? 4 : 8;
        uint32_t(&wg)[kW] = *reinterpret_cast<uint32_t(*)[kW]>(&w[SAGE_FUSE == 2 ? 0 : g * kW]);
        softmax_exp_pack16<kFp8, kM>(g, s, sc, bias, kM ? lim : 127, dsum2, wg);
        if (SAGE_FUSE == 2) {
          if constexpr (kFp8) tmem_st4(tS + g * 4, *reinterpret_cast<uint32_t(*)[4]>(&wg[0]));
          else tmem_st8(tS + g * 8, *reinterpret_cast<uint32_t(*)[8]>(&wg[0]));
        }
        if constexpr (kPs > 0) {
          if (g == kPs - 1) tmem_st_words<0, kPs * kWg>(tS, w);
          if (g == (kPs - 1 + SAGE_PSPLIT_LAG < 7 ? kPs - 1 + SAGE_PSPLIT_LAG : 7)) {
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(&p_part[t]);
          }
        }
      }
      float d0, d1;
      unpack2(dsum2, d0, d1);
      dsum = d0 + d1;
    };
    if (masked_tile) window(std::true_type{});
    else window(std::false_type{});
#else
    if (masked_tile) dsum = softmax_exp<true, 0, 128, 3, false>(s, sc, bias, lim);
    else dsum = softmax_exp<false, 0, 128, 3, false>(s, sc, bias, 127);
#endif
    }
    if (kTurn == 1 && j < n_pingpong) {
      if (t == 0) {
        named_bar_arrive(2, 2 * kTileQ);
      } else if (j < n_pingpong - 1) {
        named_bar_arrive(1, 2 * kTileQ);
      } else if (next_pingpong) {
        named_bar_arrive(1, 2 * kTileQ);
        st.granted = true;
      }
    }
    SAGE_STAMP(t, st.steps, 7);
    d_run = fmaf(d_run, alpha, dsum);
    if constexpr (kFuse == 0) softmax_pack<kFp8>(s, w);
    SAGE_STAMP(t, st.steps, 3);
    if constexpr (kPs == 0) rescale_o();
    // P row -> TMEM, over the S columns just consumed
    if (kFuse == 2) {
      // (stored group by group inside the window)
    } else if constexpr (kPs > 0) {
      tmem_st_words<kPs * kWg, 8 * kWg>(tS, w);   // (the early part went out inside the window)
    } else if constexpr (kFp8) {
      tmem_st32(tS, w);
    } else {
      uint32_t(&w0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[0]);
      uint32_t(&w1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[32]);
      tmem_st32(tS, w0);
      tmem_st32(tS + 32, w1);
    }
    tc_wait_st();
    tc_fence_before();
    if (psplit_groups<D, kFp8>() > 0 && kPs == 0) mbar_arrive(&p_part[t]);   // (attn_mask path: both parts at once)
    mbar_arrive(&p_full[t]);
    SAGE_STAMP(t, st.steps, 4);
    ++st.steps;
  }

  m_out = m_run;
  d_out = d_run;
}

// The 8 class scales of kv tile j (log2e * sm_scale * q_scale * k_scale[class]).
template <bool kVarlen>
__device__ __forceinline__ void load_class_scales(const AttnKernelParams& p, const ItemCtx& c, int j, float qs,
                                                  float (&sc)[8]) {
  const float* ksp = c.ksp;
  if constexpr (!kVarlen) {
    if (p.gran == 3) {
      // per_thread: 8 consecutive floats (2 blocks of 64 rows x 4 groups).  A second block beyond kv_len (ragged last
      // tile) has no scales: all its columns are masked, any finite factor will do - the first block's are reused
      const float4 s0 = __ldg(reinterpret_cast<const float4*>(ksp + j * 8));
      const float4 s1 = (j * 8 + 4 < p.k_scale_per_bh) ? __ldg(reinterpret_cast<const float4*>(ksp + j * 8 + 4)) : s0;
      sc[0] = s0.x * qs; sc[1] = s0.y * qs; sc[2] = s0.z * qs; sc[3] = s0.w * qs;
      sc[4] = s1.x * qs; sc[5] = s1.y * qs; sc[6] = s1.z * qs; sc[7] = s1.w * qs;
    } else {
      // per_warp / per_block: one scale per 64 kv rows
      const float a0 = __ldg(ksp + min(j * 2, p.k_scale_per_bh - 1)) * qs;
      const float a1 = __ldg(ksp + min(j * 2 + 1, p.k_scale_per_bh - 1)) * qs;
      sc[0] = sc[1] = sc[2] = sc[3] = a0;
      sc[4] = sc[5] = sc[6] = sc[7] = a1;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int idx = j * 2 + (k >> 2);             // per-block scales of the sequence, [blocks, H]
      idx = max(min(idx, c.ks_count - 1), 0);   // (0 blocks: the key-less sequence's one masked step)
      sc[k] = __ldg(ksp + (size_t)idx * c.ks_stride) * qs;
    }
  }
}

// The kv steps of one work item for Q tile `t`, upper half of S first and speculatively (SAGE_SPEC, see the top).
template <int D, bool kFp8, bool kMask, bool kVarlen = false>
__device__ __forceinline__ void softmax_steps_spec(const AttnKernelParams& p, int t, int row,
                                                   const ItemCtx& c, const WorkItem* ring, uint32_t it,
                                                   uint32_t tmem_base, uint64_t* s_full, uint64_t* s_free, uint64_t* p_full,
                                                   uint64_t* p_part, uint64_t* s_hi, WgState& st, float& m_out, float& d_out) {
  const int warp_in_wg = (row >> 5);
  const uint32_t lane_off = (uint32_t)(warp_in_wg * 32) << 16;
  const uint32_t tS = tmem_base + lane_off + (uint32_t)(t * 128);
  const uint32_t tO = tmem_base + lane_off + (uint32_t)(256 + t * D);
  const int q_row = c.q_row;
  const float qs = c.qs * p.sm_scale_log2;  // folded with sm_scale * log2 e
  const int n_iter = c.n_iter;
  constexpr int kWg = kFp8 ? 4 : 8;          // packed words per 16-column group
  constexpr float kTau = kFp8 ? kFp8Tau : kF16Tau;
  constexpr float kOff = kFp8 ? kFp8Offset : 0.f;
  float m_run = -1e30f, d_run = 0.f;
#ifdef SAGE_TRACE
  const bool trace_on = blockIdx.x == 0 && row == 0;
#endif

  for (int j = 0; j < n_iter; ++j) {
    const uint32_t ph = st.steps & 1u;
    float sc[8];
    load_class_scales<kVarlen>(p, c, j, qs, sc);
    const int kv0 = j * kTileKV;
    const int k_len = kVarlen ? c.k_len : p.Nk;
    int lim = k_len - 1 - kv0;
    if (p.causal) lim = min(lim, q_row - kv0);
    const bool masked_tile = (j == n_iter - 1) && (p.causal || (k_len - kv0) < kTileKV);
    SAGE_STAMP(t, st.steps, 0);
    uint32_t s[128];
    uint32_t w[kFp8 ? 32 : 64];
    uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
    uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
    uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
    uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
    float alpha = 1.0f, bias[8], dsum;
    uint64_t dsum2 = 0;

    auto load_hi = [&]() {
      tmem_ld32(tS + 64, s2);
      tmem_ld32(tS + 96, s3);
      tc_wait_ld();
    };
    auto wait_load_lo = [&]() {
      mbar_wait(&s_full[t], ph);
      tc_fence_after();
      SAGE_STAMP(t, st.steps, 1);
      tmem_ld32(tS + 0, s0);
      tmem_ld32(tS + 32, s1);
      tc_wait_ld();
      SAGE_STAMP(t, st.steps, 2);
    };
    auto release_s_hi = [&]() {   // the upper 64 columns of S_t are no longer needed (P only uses the lower 64)
      tc_fence_before();
      mbar_arrive(&s_free[t]);
    };
    auto debug_dump = [&]() {
      if (p.debug_s != nullptr && j == 0) {
        const int nq_pad = (p.Nq + 255) / 256 * 256;
        const int4 wv = ld_shared_v4(smem_u32(&ring[it % kSchedSlots]));
        WorkItem wi; wi.h = wv.y; wi.b = wv.z;
        int* dst = p.debug_s + ((size_t)(wi.b * p.Hq + wi.h) * nq_pad + q_row) * 128;
#pragma unroll
        for (int cc = 0; cc < 128; ++cc) dst[cc] = (int)s[cc];
      }
    };
    // exponentials + conversions of one half (4 groups of 16 columns) into its words of w
    auto exp_half = [&](auto half_c, auto masked_c) {
      constexpr int kH = decltype(half_c)::value;
      constexpr bool kM = decltype(masked_c)::value;
#pragma unroll
      for (int g = 4 * kH; g < 4 * kH + 4; ++g) {
        uint32_t(&wg)[kWg] = *reinterpret_cast<uint32_t(*)[kWg]>(&w[g * kWg]);
        softmax_exp_pack16<kFp8, kM>(g, s, sc, bias, kM ? lim : 127, dsum2, wg);
      }
    };
    auto set_bias = [&]() {
#pragma unroll
      for (int k = 0; k < 8; ++k) bias[k] = fmaf(-kMagicF, sc[k], kOff - m_run);
    };
    // rescale O when some row of this warp moved its max (PV(j-1) is complete: the lower half of S(j) was issued after
    // it and tcgen05.commit tracks all earlier MMAs) - before any part of P.V(j) may start
    auto rescale_o = [&]() {
      if (j > 0) {
        if (__any_sync(0xffffffffu, alpha != 1.0f)) {
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
      }
    };
    auto publish_hi = [&]() {   // P of keys 64..127 -> the upper half of the P columns; P.V(j) may start on them
      tmem_st_words<4 * kWg, 8 * kWg>(tS, w);
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&p_part[t]);
    };
    auto publish_lo = [&]() {
      tmem_st_words<0, 4 * kWg>(tS, w);
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&p_full[t]);
    };

    mbar_wait(&s_hi[t], ph);
    tc_fence_after();
    load_hi();
    if constexpr (kMask) {
      wait_load_lo();
      release_s_hi();
      debug_dump();
      // attn_mask: max, exponentials and row sum of y = logit + mask in one call (softmax_sm100.cuh)
      dsum = softmax_general_mask<kFp8>(s, sc, lim, c.mask_row + (size_t)kv0 * (p.mask_kind == 1 ? 1 : 2), p.mask_kind,
                                        m_run, alpha);
      softmax_pack<kFp8>(s, w);
      SAGE_STAMP(t, st.steps, 5);
      rescale_o();
      publish_hi();
      SAGE_STAMP(t, st.steps, 7);
      publish_lo();
    } else if (masked_tile) {
      wait_load_lo();
      release_s_hi();
      debug_dump();
      softmax_max<kFp8, true, false>(s, sc, lim, m_run, alpha, bias);
      rescale_o();
      exp_half(std::integral_constant<int, 1>{}, std::true_type{});
      publish_hi();
      SAGE_STAMP(t, st.steps, 5);
      exp_half(std::integral_constant<int, 0>{}, std::true_type{});
      SAGE_STAMP(t, st.steps, 7);
      publish_lo();
      float d0, d1;
      unpack2(dsum2, d0, d1);
      dsum = d0 + d1;
    } else {
      int mx[8];
      softmax_halfmax<1>(s, mx);
      float cand = -1e30f;
#pragma unroll
      for (int k = 4; k < 8; ++k) cand = fmaxf(cand, (float)mx[k] * sc[k]);
      // no row of the warp can grow beyond tau because of its upper half: start on it with the maximum it has
      const bool spec = !__any_sync(0xffffffffu, cand > m_run + kTau);
      set_bias();
      if (spec) exp_half(std::integral_constant<int, 1>{}, std::false_type{});
      SAGE_STAMP(t, st.steps, 6);
      wait_load_lo();
      debug_dump();
      softmax_halfmax<0>(s, mx);
#pragma unroll
      for (int k = 0; k < 4; ++k) cand = fmaxf(cand, (float)mx[k] * sc[k]);
      if (__any_sync(0xffffffffu, cand > m_run + kTau)) {
        const float m_new = fmaxf(m_run, cand);
        alpha = ex2_approx(m_run - m_new);  // 0 on the first tile (m_run = -1e30)
        m_run = m_new;
        set_bias();
        if (spec) {   // rare: the upper half was done against the old maximum - fetch it again and redo it
          load_hi();
          dsum2 = 0;
        }
        release_s_hi();
        exp_half(std::integral_constant<int, 1>{}, std::false_type{});
      } else {
        release_s_hi();
      }
      rescale_o();
      publish_hi();
      SAGE_STAMP(t, st.steps, 5);
      exp_half(std::integral_constant<int, 0>{}, std::false_type{});
      SAGE_STAMP(t, st.steps, 7);
      publish_lo();
      float d0, d1;
      unpack2(dsum2, d0, d1);
      dsum = d0 + d1;
    }
    d_run = fmaf(d_run, alpha, dsum);
    SAGE_STAMP(t, st.steps, 3);
    SAGE_STAMP(t, st.steps, 4);
    ++st.steps;
  }
  m_out = m_run;
  d_out = d_run;
}

// ------------------------------------------------------------------------------------------
template <int D, bool kFp8, bool kMask = false, bool kVarlen = false>
__global__ void __launch_bounds__(kThreads, 1)
sage_attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                     const AttnKernelParams p) {
  using L = SmemLayout<D, kFp8>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* q_full = bars;                    // kQStages
  uint64_t* q_empty = q_full + kQStages;      // kQStages
  uint64_t* k_full = q_empty + kQStages;      // kKStages
  uint64_t* k_empty = k_full + kKStages;      // kKStages
  uint64_t* v_full = k_empty + kKStages;      // kVStages
  uint64_t* v_empty = v_full + kVStages;      // kVStages
  uint64_t* s_full = v_empty + kVStages;      // 2
  uint64_t* p_full = s_full + 2;              // 2
  uint64_t* o_full = p_full + 2;              // 2
  uint64_t* s_free = o_full + 2;              // 2 (SAGE_SSPLIT)
  uint64_t* p_part = s_free + 2;              // 2 (SAGE_PSPLIT)
  uint64_t* s_hi = p_part + 2;                // 2 (SAGE_SSPLIT)
  uint64_t* sched_full = s_hi + 2;            // kSchedSlots
  uint64_t* sched_empty = sched_full + kSchedSlots;
  WorkItem* ring = reinterpret_cast<WorkItem*>(smem + L::kRingOff);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);

  const int warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kQStages; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], SAGE_ISSUERS); }
    for (int i = 0; i < kKStages; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], SAGE_ISSUERS); }
    for (int i = 0; i < kVStages; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], SAGE_ISSUERS); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1);
      mbar_init(&p_full[t], kTileQ);
      mbar_init(&o_full[t], 1);
      mbar_init(&s_free[t], kTileQ);
      mbar_init(&p_part[t], kTileQ);
      mbar_init(&s_hi[t], 1);
    }
    for (int i = 0; i < kSchedSlots; ++i) {
      mbar_init(&sched_full[i], 1);
      mbar_init(&sched_empty[i], kSoftmaxThreads + SAGE_ISSUERS);  // every softmax thread + the MMA issuer(s)
    }
    fence_barrier_init();
  }
  if (warp == 8) {
    if (elect_one()) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      tma_prefetch_desc(&tmO);
    }
  }
  if (warp == 9) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
#ifdef SAGE_STAGGER
  {  // experiment: take the CTAs out of lockstep
    const long long t_go = clock64() + (long long)(blockIdx.x % 8) * SAGE_STAGGER;
    while (clock64() < t_go) { }
  }
#endif

  if (warp < 8) {
    // ======================= softmax warpgroups =======================
    reg_alloc<kSoftmaxRegs>();
    const int t = warp >> 2;
    const int row = threadIdx.x & 127;
    WgState st{0u, 0u, false};
    uint32_t it = 0;
    ItemCtx c = softmax_setup<D, kVarlen>(p, t, row, ring, sched_full, it);
    while (c.valid) {
      float m_run = 0.f, d_run = 1.f;
      if (c.n_iter > 0)
#if SAGE_SPEC
        softmax_steps_spec<D, kFp8, kMask, kVarlen>(p, t, row, c, ring, it, tmem_base, s_full, s_free, p_full, p_part, s_hi,
                                                    st, m_run, d_run);
#else
        softmax_steps<D, kFp8, kMask, kVarlen>(p, t, row, c, ring, it, tmem_base, s_full, s_free, p_full, p_part, s_hi, st, m_run,
                               d_run);
#endif
      // the next item is prepared under this item's last P.V
      const bool had_steps = c.n_iter > 0;
      c = softmax_setup<D, kVarlen>(p, t, row, ring, sched_full, it + 1);
      if (had_steps) {
        softmax_epilogue<D, kFp8, kVarlen>(p, t, row, ring, it, tmem_base, smem + L::kOOff + t * L::kOBytes,
                                  reinterpret_cast<float*>(smem + L::kVsOff) + (it % kSchedSlots) * 128, &tmO, o_full, st, m_run, d_run);
      }
      mbar_arrive(&sched_empty[it % kSchedSlots]);
      ++it;
    }
    if (row == 0) tma_store_wait_read();  // the staging tile must outlive the last bulk store's reads
  } else if (warp == 8) {
    // ======================= TMA producer =======================
    reg_dealloc<kIssueRegs>();
    if (elect_one()) {
      uint32_t ip = 0, kc = 0, vc = 0;
      auto publish = [&](uint32_t idx, const WorkItem& w) {
        const uint32_t slot = idx % kSchedSlots;
        mbar_wait(&sched_empty[slot], ((idx / kSchedSlots) & 1u) ^ 1u);
        st_shared_v4(smem_u32(&ring[slot]), (uint32_t)w.q_pair, (uint32_t)w.h, (uint32_t)w.b, (uint32_t)w.hkv);
        if (p.v_scale != nullptr && w.q_pair >= 0) {
          // the item's per-channel V scales travel with the ring entry (same barrier)
          mbar_arrive_expect_tx(&sched_full[slot], D * 4);
          bulk_load_1d(smem + L::kVsOff + slot * 512, p.v_scale + (size_t)(w.b * p.Hkv + w.hkv) * D, D * 4,
                       &sched_full[slot]);
        } else {
          mbar_arrive(&sched_full[slot]);
        }
      };
      WorkItem cur = decode_work<kVarlen>(p, blockIdx.x);
      publish(0, cur);
      while (cur.q_pair >= 0) {
        // ticket of the NEXT item: issued now, consumed after this item's first loads are on their way
        const unsigned int ticket = atomicAdd(p.sched_counter, 1u);
        int n_it[2];
        item_iters_of<kVarlen>(p, cur, n_it);
        const int n_max = max(n_it[0], n_it[1]);
        const int hkv = cur.hkv;
        // kVarlen: rows are tokens of the pack (sequence offset added), the batch coordinate is 0
        const SeqInfo si = seq_info<kVarlen>(p, cur.b);
        const int cb = kVarlen ? 0 : cur.b;
        const int q_start = si.q_off + cur.q_pair * 2 * kTileQ;
        const uint32_t qb = ip % kQStages;
        mbar_wait(&q_empty[qb], ((ip / kQStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&q_full[qb], 2 * L::kQBytes);
        uint8_t* qdst = smem + L::kQOff + qb * 2 * L::kQBytes;
        tma_load_4d(qdst, &tmQ, &q_full[qb], 0, q_start, cur.h, cb);
        tma_load_4d(qdst + L::kQBytes, &tmQ, &q_full[qb], 0, q_start + kTileQ, cur.h, cb);
        WorkItem nxt;
        for (int j = 0; j < n_max; ++j) {
          const uint32_t ks = kc % kKStages, vs = vc % kVStages;
          mbar_wait(&k_empty[ks], ((kc / kKStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&k_full[ks], L::kKBytes);
          tma_load_4d(smem + L::kKOff + ks * L::kKBytes, &tmK, &k_full[ks], 0, si.k_off + j * kTileKV, hkv, cb);
          mbar_wait(&v_empty[vs], ((vc / kVStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&v_full[vs], L::kVBytes);
          uint8_t* vdst = smem + L::kVOff + vs * L::kVBytes;
          if constexpr (kFp8) {
            // V^T e4m3 [D rows][128 kv] (kv contiguous)
            tma_load_4d(vdst, &tmV, &v_full[vs], j * kTileKV, 0, hkv, cb);
          } else {
            // V f16 [128 kv][D]: one 64-wide (128 B) column block per load
#pragma unroll
            for (int c = 0; c < D / 64; ++c)
              tma_load_4d(vdst + c * (kTileKV * 128), &tmV, &v_full[vs], c * 64, si.k_off + j * kTileKV, hkv, cb);
          }
          ++kc;
          ++vc;
          if (j == 0) {
            nxt = decode_work<kVarlen>(p, gridDim.x + ticket);
            publish(ip + 1, nxt);
          }
        }
        ++ip;
        cur = nxt;
      }
    }
  } else if (warp == 9 || (SAGE_ISSUERS == 2 && warp == 10)) {
    // ======================= MMA issuer(s) =======================
    reg_dealloc<kIssueRegs>();
    // the Q tiles this warp issues for: both, or one each
    const int t_lo = (SAGE_ISSUERS == 2) ? warp - 9 : 0, t_hi = (SAGE_ISSUERS == 2) ? warp - 8 : 2;
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

      uint32_t im = 0, kc = 0, vc = 0;  // items started, K tiles consumed, V tiles consumed
      uint32_t np[2] = {0, 0};          // PV_t issued so far

      // S_t = Q_t.K^T of the K tile in ring slot (kc % stages).  No "S drained" barrier is needed: the
      // issuer only gets here after it has issued the P.V that reads P_t (which the warpgroup wrote
      // after pulling S_t into registers), and the tensor pipe executes in order.
      auto issue_qk = [&](int t, uint32_t qb) {
        const uint32_t a0 = sQ + (qb * 2 + t) * L::kQBytes;
        const uint32_t b0 = sK + (kc % kKStages) * L::kKBytes;
#pragma unroll
        for (int kk = 0; kk < D / 32; ++kk) {
          const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, qk_sbo, qk_swz);
          const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, qk_sbo, qk_swz);
          mma_i8_ss(tmem_base + t * 128, ad, bd, idesc_qk, kk > 0);
        }
        tc_commit(&s_full[t]);
      };
      // SAGE_SSPLIT: the two N = 64 halves of S_t (hi: keys 64..127 -> columns [64, 128); lo: keys 0..63)
      constexpr uint32_t idesc_qk_half = make_idesc(CFMT_S32, FMT_S8, FMT_S8, 0, 0, kTileQ, kTileKV / 2);
      uint32_t ns[2] = {0, 0};          // S_t issued so far (whole or split)
      auto issue_qk_half = [&](int t, uint32_t qb, int hi) {
        const uint32_t a0 = sQ + (qb * 2 + t) * L::kQBytes;
        const uint32_t b0 = sK + (kc % kKStages) * L::kKBytes + (hi ? (kTileKV / 2) * D : 0);
#pragma unroll
        for (int kk = 0; kk < D / 32; ++kk) {
          const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, qk_sbo, qk_swz);
          const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, qk_sbo, qk_swz);
          mma_i8_ss(tmem_base + t * 128 + (hi ? 64 : 0), ad, bd, idesc_qk_half, kk > 0);
        }
      };
      auto issue_pv = [&](int t, bool accumulate) {
        const uint32_t b0 = sV + (vc % kVStages) * L::kVBytes;
        const uint32_t d_tmem = tmem_base + 256 + t * D;
        const uint32_t a_tmem = tmem_base + t * 128;
        constexpr int kSteps = kFp8 ? kTileKV / 32 : kTileKV / 16;
        constexpr int kPsplit = SAGE_SPEC ? 4 : psplit_groups<D, kFp8>();
        constexpr int kEarly = kFp8 ? kPsplit / 2 : kPsplit;   // K-steps covered by the early part of P
        if (kEarly > 0) {
          mbar_wait(&p_part[t], np[t] & 1u);
          tc_fence_after();
        }
#pragma unroll
        for (int ki = 0; ki < kSteps; ++ki) {
          if (ki == kEarly) {
            mbar_wait(&p_full[t], np[t] & 1u);
            tc_fence_after();
          }
          // SAGE_SPEC: the early part of P is the UPPER half of the keys
          const int kk = SAGE_SPEC ? (ki + kSteps / 2) % kSteps : ki;
          const bool acc = accumulate || (ki > 0);
          if constexpr (kFp8) {
            const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, 1024, SWZ_128B);
            mma_f8_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, acc);
          } else {
            // MN-major B: 16 kv rows x 128 B per MMA = 2 KB step; 64-wide d blocks 16 KB apart
            const uint64_t bd = make_smem_desc(b0 + kk * 2048, kTileKV * 128, 1024, SWZ_128B);
            mma_f16_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, acc);
          }
        }
        ++np[t];
      };

      int n_it[2], n_nx[2];
      bool have = false;
      {
        const WorkItem w0 = read_work(ring, sched_full, 0);
        mbar_arrive(&sched_empty[0]);
        have = w0.q_pair >= 0;
        if (have) {
          item_iters_of<kVarlen>(p, w0, n_it);
          mbar_wait(&q_full[0], 0);
          mbar_wait(&k_full[0], 0);
          tc_fence_after();
          for (int t = t_lo; t < t_hi; ++t)
            if (n_it[t] > 0) {
              if (SAGE_SSPLIT) {
                issue_qk_half(t, 0, 1);
                tc_commit(&s_hi[t]);
                issue_qk_half(t, 0, 0);
                tc_commit(&s_full[t]);
              } else {
                issue_qk(t, 0);
              }
              ++ns[t];
            }
          tc_commit(&k_empty[0]);
          ++kc;
          if (max(n_it[0], n_it[1]) == 1) tc_commit(&q_empty[0]);
        }
      }
      while (have) {
        const WorkItem wn = read_work(ring, sched_full, im + 1);
        mbar_arrive(&sched_empty[(im + 1) % kSchedSlots]);
        const bool have_nxt = wn.q_pair >= 0;
        if (have_nxt) item_iters_of<kVarlen>(p, wn, n_nx);
        const int n_max = max(n_it[0], n_it[1]);
        const uint32_t qb = im % kQStages;
        for (int j = 0; j < n_max; ++j) {
          mbar_wait(&v_full[vc % kVStages], (vc / kVStages) & 1u);
          // what follows P_t.V(j) on the tensor pipe: S_t(j+1) of this item, or S_t(0) of the next
          const bool more = (j + 1 < n_max);
          const bool cross = !more && have_nxt;
          const int jn = more ? j + 1 : 0;
          const int nn0 = more ? n_it[0] : n_nx[0], nn1 = more ? n_it[1] : n_nx[1];
          const uint32_t qn = more ? qb : (im + 1) % kQStages;
          if (more || cross) {
            if (cross) mbar_wait(&q_full[qn], ((im + 1) / kQStages) & 1u);
            mbar_wait(&k_full[kc % kKStages], (kc / kKStages) & 1u);
            tc_fence_after();
          }
          for (int t = t_lo; t < t_hi; ++t) {
            if (SAGE_SSPLIT && (more || cross) && jn < (t == 0 ? nn0 : nn1)) {
              // upper half of the next S_t first: it only needs the warpgroup to have drained the current one
              if (ns[t] > 0) mbar_wait(&s_free[t], (ns[t] - 1) & 1u);
              tc_fence_after();
              issue_qk_half(t, qn, 1);
              tc_commit(&s_hi[t]);
            }
            if (j < n_it[t]) {
              issue_pv(t, j > 0);
              if (j == n_it[t] - 1) tc_commit(&o_full[t]);
            }
            if (t == t_hi - 1) tc_commit(&v_empty[vc % kVStages]);
            if ((more || cross) && jn < (t == 0 ? nn0 : nn1)) {
              if (SAGE_SSPLIT) {
                issue_qk_half(t, qn, 0);
                tc_commit(&s_full[t]);
              } else {
                issue_qk(t, qn);
              }
              ++ns[t];
            }
          }
          ++vc;
          if (more || cross) {
            tc_commit(&k_empty[kc % kKStages]);
            ++kc;
            if (jn == max(nn0, nn1) - 1) tc_commit(&q_empty[qn]);  // last use of that Q slot
          }
        }
        ++im;
        n_it[0] = n_nx[0];
        n_it[1] = n_nx[1];
        have = have_nxt;
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

}  // namespace sage_attn

// ==========================================================================================
// Host side: TMA descriptors + launch
// ==========================================================================================
namespace sage_attn {
using namespace sage;

template <int D, bool kFp8, bool kMask = false, bool kVarlen = false>
static int launch_attn(const sage_attn_args* a, cudaStream_t stream) {
  using L = SmemLayout<D, kFp8>;
  CUtensorMap tmQ, tmK, tmV, tmO;
  // real head_dim of o (and of an f16 v): columns [d_og, D) are zero-filled on load and clipped on store
  const int d_og = a->d_og > 0 ? a->d_og : D;
  SAGE_REQUIRE(d_og <= D && d_og % 8 == 0, "d_og (%d) must be a multiple of 8 and <= head_dim %d", d_og, D);
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
    rc = make_map4(&tmV, CU_TENSOR_MAP_DATA_TYPE_UINT16, a->v.ptr, d_og, a->Nk, a->Hkv, a->B,
                   a->v.stride_n * 2, a->v.stride_h * 2, a->v.stride_b * 2, 64, kTileKV,
                   CU_TENSOR_MAP_SWIZZLE_128B, "v_f16");
  }
  if (rc) return rc;
  // o rows through TMA: 64-column (128 B) blocks of 128 rows; rows >= Nq are clipped by the hardware
  rc = make_map4(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, a->o.ptr, d_og, a->Nq, a->Hq, a->B,
                 a->o.stride_n * 2, a->o.stride_h * 2, a->o.stride_b * 2, 64, kTileQ,
                 CU_TENSOR_MAP_SWIZZLE_128B, "o");
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

  SAGE_REQUIRE(kMask == (a->mask != nullptr), "internal: mask kernel selection");
  p.mask = reinterpret_cast<const uint8_t*>(a->mask);
  p.mask_kind = a->mask_kind;
  p.mask_sb = a->mask_stride_b; p.mask_sh = a->mask_stride_h; p.mask_sm = a->mask_stride_m;
  SAGE_REQUIRE(kVarlen == (a->cu_seqlens_q != nullptr), "internal: varlen kernel selection");
  p.cu_q = a->cu_seqlens_q; p.cu_k = a->cu_seqlens_k; p.cu_q_scale = a->cu_seqlens_q_scale;
  p.cu_k_scale = a->cu_seqlens_k_scale; p.cu_pairs = a->cu_pairs; p.nseq = a->nseq;
  p.o_peer_maps = nullptr;
  p.o_peer_ptrs = nullptr;
  p.o_peer_stride_h = p.o_peer_stride_n = 0;
  p.o_peer_rows = p.o_peer_count = 0;
  if (a->o_peer_ptrs != nullptr) {
    SAGE_REQUIRE(!kVarlen && a->B == 1, "o_peer_ptrs: fixed-length calls with B = 1 only");
    SAGE_REQUIRE(a->o_peer_count > 0 && a->o_peer_count <= 16 && a->o_peer_rows > 0 && a->o_peer_maps != nullptr &&
                 (reinterpret_cast<uintptr_t>(a->o_peer_maps) & 63) == 0,
                 "o_peer_*: 1..16 destinations, rows per destination > 0, a 64-byte aligned device scratch of SAGE_O_PEER_SCRATCH_BYTES");
    SAGE_REQUIRE((long long)a->o_peer_count * a->o_peer_rows >= a->Nq, "o_peer_count * o_peer_rows (%d * %d) < Nq (%d)",
                 a->o_peer_count, a->o_peer_rows, a->Nq);
    // scratch layout: the tensor maps (128 bytes each), then the base addresses
    struct alignas(64) { CUtensorMap maps[16]; void* ptrs[16]; } host;
    CUtensorMap* host_maps = host.maps;
    for (int s_ = 0; s_ < a->o_peer_count; ++s_) {
      host.ptrs[s_] = a->o_peer_ptrs[s_];
      rc = make_map4(&host_maps[s_], CU_TENSOR_MAP_DATA_TYPE_UINT16, a->o_peer_ptrs[s_], d_og, a->o_peer_rows, a->Hq, 1,
                     a->o_peer_stride_n * 2, a->o_peer_stride_h * 2,
                     (uint64_t)a->o_peer_stride_n * 2 * (uint64_t)a->o_peer_rows, 64, kTileQ, CU_TENSOR_MAP_SWIZZLE_128B,
                     "o_peer");
      if (rc) return rc;
    }
    // (pageable source: staged before the call returns)
    SAGE_CUDA(cudaMemcpyAsync(a->o_peer_maps, host_maps, sizeof(CUtensorMap) * a->o_peer_count, cudaMemcpyHostToDevice,
                              stream));
    void* dev_ptrs = reinterpret_cast<uint8_t*>(a->o_peer_maps) + sizeof(CUtensorMap) * 16;
    SAGE_CUDA(cudaMemcpyAsync(dev_ptrs, host.ptrs, sizeof(void*) * a->o_peer_count, cudaMemcpyHostToDevice, stream));
    p.o_peer_maps = reinterpret_cast<const CUtensorMap*>(a->o_peer_maps);
    p.o_peer_ptrs = reinterpret_cast<void* const*>(dev_ptrs);
    p.o_peer_stride_h = a->o_peer_stride_h;
    p.o_peer_stride_n = a->o_peer_stride_n;
    p.o_peer_rows = a->o_peer_rows;
    p.o_peer_count = a->o_peer_count;
  }
  auto kern = sage_attn_fwd_kernel<D, kFp8, kMask, kVarlen>;
  // per template instantiation AND per device: function attributes belong to the device's context
  static SmemAttrOnce smem_attr;
  SAGE_CUDA(smem_attr.ensure(kern, L::kDynamic));
  p.n_qpairs = (a->Nq + 2 * kTileQ - 1) / (2 * kTileQ);
  const long long n_work = (long long)a->B * a->Hq * p.n_qpairs;
  SAGE_REQUIRE(n_work < (1ll << 31), "too many work items");
  p.n_work = (int)n_work;
  p.sched_counter = acquire_sched_counter(stream);
  if (p.sched_counter == nullptr) return -1;
  SAGE_CUDA(cudaMemsetAsync(p.sched_counter, 0, sizeof(unsigned int), stream));
  int ctas = sm_count();
  if (a->max_ctas > 0 && a->max_ctas < ctas) ctas = a->max_ctas;
  // kVarlen: the number of items is only known on the device (cu_pairs): one CTA per SM, the surplus ones leave at once
  const int grid = kVarlen ? ctas : (int)(n_work < ctas ? n_work : ctas);
  kern<<<grid, kThreads, L::kDynamic, stream>>>(tmQ, tmK, tmV, tmO, p);
  SAGE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace sage_attn

