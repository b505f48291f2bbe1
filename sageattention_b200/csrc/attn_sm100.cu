// SageAttention forward for sm_100a (B200), FP8 P.V flavour: INT8 QK^T + online softmax + e4m3 P.V
// (sage_attn_i8_f8; the f16 P.V entry point lives in attn_f16_sm100.cu)
//
// Replaces the reference's mma.sync kernels
//   qk_int_sv_f16_attn_kernel  (/root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:44-671)
//   qk_int_sv_f8_attn_kernel   (/root/reference/csrc/qattn/qk_int_sv_f8_cuda_sm89.cuh:44-707)
// with a from-scratch tcgen05 design (not a port):
//
//   * PERSISTENT grid (one CTA per SM): a work item is TWO 128-row Q tiles (256 query rows) of one
//     (batch, head), walked over the KV sequence in 128-row tiles; items come from a dynamic scheduler
//     (attn_common_sm100.cuh), barrier phases run on across items, and the issuer launches the next
//     item's first Q.K^T under the current item's last step and epilogue;
//   * warp 8 is the TMA producer (a Q ring of two items, a K ring and a V ring, 128B-swizzled smem);
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
#include <atomic>
#include <type_traits>

#include "attn_common_sm100.cuh"

namespace sage {

// 1 = the exp phases of the two softmax warpgroups alternate (named barriers), 0 = free-running.
// (Whole-step alternation, a fixed phase offset and pre-biased S were measured slower and live in
// tools/experiments/attn_v0e_f8_per_item_grid.cu.)
#ifndef SAGE_TURN
#define SAGE_TURN 1
#endif
// 1: S(j+1) is pulled into registers under the P store's fence of step j instead of at the top of step
// j+1.  Measured 6 % SLOWER (hd128 1288 -> 1211, hd64 674 -> 623 TOPS at N=8192), default 0.
#ifndef SAGE_SPREF
#define SAGE_SPREF 0
#endif
// k scales of later steps prefetched to L1: measured -0.7 %, off.
#ifndef SAGE_KPREF
#define SAGE_KPREF 0
#endif

// 1: the unmasked row maximum as depth-3 trees of 3-input maxima (40 independent first-level operations)
#ifndef SAGE_MAXTREE
#define SAGE_MAXTREE 0
#endif
// 1: the k scales of a kv step as two 16-byte loads (per_thread) / two scalars (per_warp) instead of eight indexed loads.
// Measured (one box, kernel-only, N = 8K / 32K): +3 % in the f16 hd128 kernel (attn_f16_sm100.cu: 1253 -> 1299 / 1161 ->
// 1198 TOPS, where it is on) but -2 % (fp8 hd128: 1304 -> 1283) and -5 % (fp8 hd64: 679 -> 648) here, so off.
#ifndef SAGE_KSVEC
#define SAGE_KSVEC 0
#endif
// 0: exp, pack and store are separate phases.  1: the e4m3 / f16 conversions are issued group by group (16
// columns) inside the exp phase, under the MUFU pipe.  2: as 1, and every group is stored (shared memory or
// TMEM) as soon as it is packed, so that only the fence and the barrier arrival are left after the window.
#ifndef SAGE_FUSE
#define SAGE_FUSE 0
#endif

// P_t in its own TMEM columns (head_dim 64: S 256 + O 128 + P 64/128 columns) or in shared memory
template <int D>
constexpr bool kPTmem = (D == 64);

template <int D, bool kFp8>
struct SmemLayout {
  static constexpr int kQBytes = kTileQ * D;                            // int8
  static constexpr int kKBytes = kTileKV * D;                           // int8
  static constexpr int kVBytes = kFp8 ? D * kTileKV : kTileKV * D * 2;  // e4m3 [D][128] | f16 [128][D]
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + kQStages * 2 * kQBytes;
  static constexpr int kVOff = kKOff + kKStages * kKBytes;
  static constexpr int kPBytes = kPTmem<D> ? 0 : (kFp8 ? kTileQ * kTileKV : kTileQ * kTileKV * 2);  // K-major, SW128
  static constexpr int kPOff = kVOff + kVStages * kVBytes;
  static constexpr int kOBytes = kTileQ * D * 2;  // output staging tile per warpgroup: D/64 blocks of [128][128 B], SW128
  static constexpr int kOOff = kPOff + 2 * kPBytes;
  static constexpr int kBarOff = kOOff + 2 * kOBytes;
  static constexpr int kNumBars = 2 * kQStages + 2 * kKStages + 2 * kVStages + 5 * 2 + 2 * 3 /*work ring*/;
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
                                              uint32_t tmem_base, uint8_t* p_smem, uint64_t* s_full,
                                              uint64_t* s_free, uint64_t* p_full, uint64_t* p_empty,
                                              WgState& st, float& m_out, float& d_out) {
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
  // (the attn_mask kernels run their warpgroups freely: kTurn = 0)
  constexpr int kTurn = kMask ? 0 : SAGE_TURN;
  if (kTurn != 0 && t == 1 && n_pingpong > 0) {
    if (!st.granted) named_bar_arrive(1, 2 * kTileQ);
    st.granted = false;
  }
#ifdef SAGE_TRACE
  const bool trace_on = blockIdx.x == 0 && row == 0;
#endif

  uint32_t s[128];
  auto load_s = [&]() {
    uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
    uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
    uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
    uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
    tmem_ld32(tS + 0, s0);
    tmem_ld32(tS + 32, s1);
    tmem_ld32(tS + 64, s2);
    tmem_ld32(tS + 96, s3);
  };
#if SAGE_SPREF
  // the first S of the item; every later one is pulled in under the previous step's P store
  mbar_wait(&s_full[t], st.steps & 1u);
  tc_fence_after();
  load_s();
  tc_wait_ld();
  tc_fence_before();
  mbar_arrive(&s_free[t]);  // S_t is in registers: the issuer may overwrite it with the next S_t
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
    // pull the next steps' k scales towards L1 (a 128-byte line holds 4 (per_thread) / 16 steps)
    if (SAGE_KPREF) prefetch_l1(ksp + min((j + 2) * (p.gran == 3 ? 8 : 2), p.k_scale_per_bh - 1));
    SAGE_STAMP(t, st.steps, 0);
#if !SAGE_SPREF
    mbar_wait(&s_full[t], ph);
    tc_fence_after();
    SAGE_STAMP(t, st.steps, 1);
    load_s();
    tc_wait_ld();
    // S_t is in registers: the issuer may overwrite it with the next S_t
    tc_fence_before();
    mbar_arrive(&s_free[t]);
#endif
    SAGE_STAMP(t, st.steps, 2);
    if (p.debug_s != nullptr && j == 0) {
      const int nq_pad = (p.Nq + 255) / 256 * 256;
      const int4 wv = ld_shared_v4(smem_u32(&ring[it % kSchedSlots]));
      WorkItem wi; wi.h = wv.y; wi.b = wv.z;
      int* dst = p.debug_s + ((size_t)(wi.b * p.Hq + wi.h) * nq_pad + q_row) * 128;
#pragma unroll
      for (int c = 0; c < 128; ++c) dst[c] = (int)s[c];
    }
    const int kv0 = j * kTileKV;
    const int k_len = kVarlen ? c.k_len : p.Nk;
    int lim = k_len - 1 - kv0;
    if (p.causal) lim = min(lim, q_row - kv0);
    const bool masked_tile = (j == n_iter - 1) && (p.causal || (k_len - kv0) < kTileKV);
    float alpha, bias[8], dsum;
    if constexpr (kMask) {
      // attn_mask: max, exponentials and row sum of y = logit + mask in one call (softmax_sm100.cuh)
      dsum = softmax_general_mask<kFp8>(s, sc, lim, c.mask_row + (size_t)kv0 * (p.mask_kind == 1 ? 1 : 2), p.mask_kind,
                                        m_run, alpha);
    } else if (masked_tile) softmax_max<kFp8, true, false>(s, sc, lim, m_run, alpha, bias);
    else if (SAGE_MAXTREE) softmax_max_tree<kFp8>(s, sc, m_run, alpha, bias);
    else softmax_max<kFp8, false, false>(s, sc, 127, m_run, alpha, bias);
    SAGE_STAMP(t, st.steps, 5);
    // shared-memory address of this row's P (K-major, 128B swizzle: 16B chunk c of row r sits at chunk c ^ (r & 7))
    const uint32_t prow = smem_u32(p_smem) + (uint32_t)row * 128u;
    const uint32_t px = (uint32_t)(row & 7);
    const uint32_t tP = tmem_base + lane_off + (uint32_t)(384 + t * (kFp8 ? 32 : 64));  // kPTmem only
#if SAGE_FUSE
    // PV(j-1) complete: P tile reusable (checked before the window when the groups are stored inside it)
    if (SAGE_FUSE == 2 && j > 0) mbar_wait(&p_empty[t], ph ^ 1u);
    if (kTurn == 1 && j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
    SAGE_STAMP(t, st.steps, 6);
    uint32_t w[kFp8 ? 32 : 64];  // (SAGE_FUSE == 2 reuses the first group's words)
    auto window = [&](auto masked_c) {
      constexpr bool kM = decltype(masked_c)::value;
      uint64_t dsum2 = 0;
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        constexpr int kW = kFp8 ? 4 : 8;
        uint32_t(&wg)[kW] = *reinterpret_cast<uint32_t(*)[kW]>(&w[SAGE_FUSE == 2 ? 0 : g * kW]);
        softmax_exp_pack16<kFp8, kM>(g, s, sc, bias, kM ? lim : 127, dsum2, wg);
        if (SAGE_FUSE == 2) {
          if constexpr (kPTmem<D>) {
            if constexpr (kFp8) tmem_st4(tP + g * 4, *reinterpret_cast<uint32_t(*)[4]>(&wg[0]));
            else tmem_st8(tP + g * 8, *reinterpret_cast<uint32_t(*)[8]>(&wg[0]));
          } else {
            // (this file's shared-memory P is e4m3: 16 columns = one 16-byte chunk)
            st_shared_v4(prow + (((uint32_t)g ^ px) << 4), wg[0], wg[1], wg[2], wg[3]);
          }
        }
      }
      float d0, d1;
      unpack2(dsum2, d0, d1);
      dsum = d0 + d1;
    };
    if (masked_tile) window(std::true_type{});
    else window(std::false_type{});
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
    SAGE_STAMP(t, st.steps, 3);
    if (SAGE_FUSE != 2 && j > 0) mbar_wait(&p_empty[t], ph ^ 1u);
#else
    uint32_t w[kFp8 ? 32 : 64];
    // The exp phases of the two warpgroups alternate (they are the only MUFU-bound stretch): each
    // warpgroup's load / max / pack / store phases then run under the other one's exponentials.
    if (kTurn == 1 && j < n_pingpong) named_bar_sync(1 + t, 2 * kTileQ);
    SAGE_STAMP(t, st.steps, 6);
    if constexpr (kMask) { /* done above */ }
    else if (masked_tile) dsum = softmax_exp<true, 0, 128, 3, false>(s, sc, bias, lim);
    else dsum = softmax_exp<false, 0, 128, 3, false>(s, sc, bias, 127);
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
    softmax_pack<kFp8>(s, w);
    SAGE_STAMP(t, st.steps, 3);
    // PV(j-1) complete: P tile reusable and O_t stable (at j == 0 the epilogue of the previous item
    // has already waited for that item's last PV)
    if (j > 0) mbar_wait(&p_empty[t], ph ^ 1u);
#endif
    // P leaves the registers first (w[] is dead after this) ...
    if (SAGE_FUSE != 2) {
      if constexpr (kPTmem<D>) {
        // head_dim 64 leaves 128 TMEM columns free: P_t gets its own columns there (no shared-memory
        // store, no proxy fence, and the P.V MMA reads A from TMEM instead of the shared-memory port)
        if constexpr (kFp8) {
          tmem_st32(tP, *reinterpret_cast<uint32_t(*)[32]>(&w[0]));
        } else {
          tmem_st32(tP, *reinterpret_cast<uint32_t(*)[32]>(&w[0]));
          tmem_st32(tP + 32, *reinterpret_cast<uint32_t(*)[32]>(&w[kFp8 ? 0 : 32]));
        }
      } else {
#pragma unroll
        for (int c = 0; c < (kFp8 ? 8 : 16); ++c) {
          const uint32_t addr = prow + (uint32_t)(c >> 3) * (kTileQ * 128) + (((uint32_t)(c & 7) ^ px) << 4);
          st_shared_v4(addr, w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
        }
      }
    }
    // ... then O is rescaled when some row of this warp moved its max ...
    if (j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {
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
    }
    // ... and the logits of the NEXT step (ready long ago: their MMA went out when this step's S was
    // drained) start their way into the s[] registers, under the fences and the barrier traffic.
    const bool s_pref = SAGE_SPREF && (j + 1 < n_iter);
    if (s_pref) {
      mbar_wait(&s_full[t], ph ^ 1u);
      tc_fence_after();
      load_s();
    }
    if constexpr (kPTmem<D>) {
      tc_wait_st();
    } else {
      tc_wait_st();  // the rescale's stores (no-op otherwise)
      fence_proxy_async_smem();
    }
    tc_fence_before();
    mbar_arrive(&p_full[t]);
    if (s_pref) {
      tc_wait_ld();
      tc_fence_before();
      mbar_arrive(&s_free[t]);
    }
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
  uint64_t* s_free = s_full + 2;              // 2
  uint64_t* p_full = s_free + 2;              // 2
  uint64_t* p_empty = p_full + 2;             // 2
  uint64_t* o_full = p_empty + 2;             // 2
  uint64_t* sched_full = o_full + 2;          // kSchedSlots
  uint64_t* sched_empty = sched_full + kSchedSlots;
  WorkItem* ring = reinterpret_cast<WorkItem*>(smem + L::kRingOff);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);

  const int warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kQStages; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < kKStages; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < kVStages; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1);
      mbar_init(&s_free[t], kTileQ);
      mbar_init(&p_full[t], kTileQ);
      mbar_init(&p_empty[t], 1);
      mbar_init(&o_full[t], 1);
    }
    for (int i = 0; i < kSchedSlots; ++i) {
      mbar_init(&sched_full[i], 1);
      mbar_init(&sched_empty[i], kSoftmaxThreads + 1);  // every softmax thread + the MMA issuer
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
        softmax_steps<D, kFp8, kMask, kVarlen>(p, t, row, c, ring, it, tmem_base, smem + L::kPOff + t * L::kPBytes,
                               s_full, s_free, p_full, p_empty, st, m_run, d_run);
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
      const uint32_t sP = smem_u32(smem + L::kPOff);

      uint32_t im = 0, kc = 0, vc = 0;       // items started, K tiles consumed, V tiles consumed
      uint32_t ns[2] = {0, 0}, np[2] = {0, 0};  // S_t / PV_t issued so far

      // S_t = Q_t.K^T of the K tile in ring slot (kc % stages); waits for the warpgroup to have drained
      // the previous S_t
      auto issue_qk = [&](int t, uint32_t qb) {
        if (ns[t] > 0) mbar_wait(&s_free[t], (ns[t] - 1) & 1u);
        tc_fence_after();
        const uint32_t a0 = sQ + (qb * 2 + t) * L::kQBytes;
        const uint32_t b0 = sK + (kc % kKStages) * L::kKBytes;
#pragma unroll
        for (int kk = 0; kk < D / 32; ++kk) {
          const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, qk_sbo, qk_swz);
          const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, qk_sbo, qk_swz);
          mma_i8_ss(tmem_base + t * 128, ad, bd, idesc_qk, kk > 0);
        }
        tc_commit(&s_full[t]);
        ++ns[t];
      };
#ifdef SAGE_TRACE
      const bool trace_on = blockIdx.x == 0;
#endif
      auto issue_pv = [&](int t, bool accumulate) {
        SAGE_STAMP(2, np[t], t * 4 + 0);
        mbar_wait(&p_full[t], np[t] & 1u);
        tc_fence_after();
        SAGE_STAMP(2, np[t], t * 4 + 1);
        const uint32_t a0 = sP + t * L::kPBytes;
        const uint32_t b0 = sV + (vc % kVStages) * L::kVBytes;
        const uint32_t d_tmem = tmem_base + 256 + t * D;
        const uint32_t a_tmem = tmem_base + 384 + t * (kFp8 ? 32 : 64);  // kPTmem only
        if constexpr (kFp8) {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 32; ++kk) {
            const uint64_t bd = make_smem_desc(b0 + kk * 32, 16, 1024, SWZ_128B);
            if constexpr (kPTmem<D>) {
              mma_f8_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, accumulate || (kk > 0));
            } else {
              const uint64_t ad = make_smem_desc(a0 + kk * 32, 16, 1024, SWZ_128B);
              mma_f8_ss(d_tmem, ad, bd, idesc_pv, accumulate || (kk > 0));
            }
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < kTileKV / 16; ++kk) {
            // MN-major B: 16 kv rows x 128 B per MMA = 2 KB step; 64-wide d blocks 16 KB apart
            const uint64_t bd = make_smem_desc(b0 + kk * 2048, kTileKV * 128, 1024, SWZ_128B);
            if constexpr (kPTmem<D>) {
              mma_f16_ts(d_tmem, a_tmem + kk * 8, bd, idesc_pv, accumulate || (kk > 0));
            } else {
              // A: two 64-wide K blocks of [128 rows][128 B], 16 KB apart, 32 B per MMA inside a block
              const uint64_t ad = make_smem_desc(a0 + (kk >> 2) * (kTileQ * 128) + (kk & 3) * 32, 16, 1024, SWZ_128B);
              mma_f16_ss(d_tmem, ad, bd, idesc_pv, accumulate || (kk > 0));
            }
          }
        }
        tc_commit(&p_empty[t]);
        SAGE_STAMP(2, np[t], t * 4 + 2);
        ++np[t];
      };
      // Q.K^T of kv step `jj` of the item in Q slot `qb` (both tiles), K from the ring
      auto qk_step = [&](const int (&nit)[2], int jj, uint32_t qb) {
        mbar_wait(&k_full[kc % kKStages], (kc / kKStages) & 1u);
#pragma unroll
        for (int t = 0; t < 2; ++t)
          if (jj < nit[t]) issue_qk(t, qb);
        tc_commit(&k_empty[kc % kKStages]);
        ++kc;
        if (jj == max(nit[0], nit[1]) - 1) tc_commit(&q_empty[qb]);  // last use of this Q slot
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
          qk_step(n_it, 0, 0);
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
          // the next Q.K^T (of this item, or the first one of the next item) goes out before the
          // issuer blocks on this step's P tiles
          if (j + 1 < n_max) {
            qk_step(n_it, j + 1, qb);
          } else if (have_nxt) {
            const uint32_t qn = (im + 1) % kQStages;
            mbar_wait(&q_full[qn], ((im + 1) / kQStages) & 1u);
            qk_step(n_nx, 0, qn);
          }
          mbar_wait(&v_full[vc % kVStages], (vc / kVStages) & 1u);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (j < n_it[t]) {
              issue_pv(t, j > 0);
              if (j == n_it[t] - 1) tc_commit(&o_full[t]);
            }
          }
          tc_commit(&v_empty[vc % kVStages]);
          ++vc;
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

}  // namespace sage

// ==========================================================================================
// Host side: TMA descriptors + launch
// ==========================================================================================
namespace sage {

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

}  // namespace sage

// head_dim 64 with f16 P also runs on this kernel (P in spare TMEM columns, S(j+1) issued as soon as
// S(j) is drained); called from sage_attn_i8_f16 in attn_f16_sm100.cu
namespace sage {
int launch_attn_f16_d64(const sage_attn_args* a, cudaStream_t stream) {
  if (a->cu_seqlens_q != nullptr) return launch_attn<64, false, false, true>(a, stream);
  return a->mask ? launch_attn<64, false, true>(a, stream) : launch_attn<64, false>(a, stream);
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
  SAGE_REQUIRE(a->cu_seqlens_q == nullptr, "varlen is an f16-P.V path (sage_attn_i8_f16)");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (a->mask != nullptr)
    return a->D == 128 ? sage::launch_attn<128, true, true>(a, s) : sage::launch_attn<64, true, true>(a, s);
  return a->D == 128 ? sage::launch_attn<128, true>(a, s) : sage::launch_attn<64, true>(a, s);
}
