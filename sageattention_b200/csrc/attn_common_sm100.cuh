// Support code of the attention kernel (attn_kernel_sm100.cuh, instantiated by attn_f16_sm100.cu and attn_f8_sm100.cu):
// kernel parameters, the persistent work list with its dynamic scheduler, the per-item set-up and epilogue of the softmax
// warpgroups, and the host-side helpers (TMA descriptors, per-device state, argument checks).  The step loop, the MMA
// issuer, the kernel and its launcher are in attn_kernel_sm100.cuh.  Everything here has internal linkage or is a
// template / inline.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <map>
#include <mutex>
#include <thread>
#include <cstdio>
#include <cstring>

#include "../../include/sage_b200.h"
#include "ptx_sm100.cuh"
#include "softmax_sm100.cuh"
#include "status.h"

namespace sage {

// Development aid (-DSAGE_TRACE, tools/trace_attn.py): SM-clock stamps of CTA 0's first kv steps
// (counted across its work items; slots 8/9 = last P.V seen / end of the item's epilogue).
#ifdef SAGE_TRACE
constexpr int kTraceSteps = 64, kTraceSlots = 12;
static __device__ long long g_trace[3][kTraceSteps][kTraceSlots];  // one copy per translation unit
#define SAGE_STAMP(who, j, slot)                                                             \
  do {                                                                                       \
    if (trace_on && (j) < kTraceSteps) g_trace[who][j][slot] = clock64();                    \
  } while (0)
#else
#define SAGE_STAMP(who, j, slot) do { } while (0)
#endif


constexpr int kTileQ = 128;       // rows per Q tile (= TMEM lanes)
constexpr int kTileKV = 128;      // kv rows per step
constexpr int kQStages = 2;      // Q tiles of the current and of the next work item
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
  int n_qpairs;  // ceil(Nq / 256)
  int n_work;    // work items: B * Hq * n_qpairs
  unsigned int* sched_counter;  // zeroed on the launch stream just before the kernel; see g_sched_counter
  int out_bf16;
  float sm_scale_log2;
  // attn_mask (kMask kernels only): [B|1, H|1, Nq, >= ceil128(Nk)] with broadcast dims at stride 0, rows 16-byte
  // aligned and readable up to the end of the last kv tile; strides in elements
  const uint8_t* mask;
  int mask_kind;  // 0 none, 1 bool (u8, non-zero = attend), 2 f16 additive, 3 bf16 additive
  long long mask_sb, mask_sh, mask_sm;
  // packed variable-length batches (kVarlen kernels, sageattn_varlen): q / k / v / o are [tokens, H, D] (B = 1 in the
  // tensor maps), a work item's batch index is the SEQUENCE, scales are [blocks, H] with sequence s starting at row
  // cu_*_scale[s], cu_pairs = prefix of ceil(q_len / 256) (all written by sage_quant_per_block_varlen_int8)
  const int* cu_q; const int* cu_k; const int* cu_q_scale; const int* cu_k_scale; const int* cu_pairs;
  int nseq;
  // sequence-parallel output scatter (sage_attn_args.o_peer_*): rows [s * o_peer_rows, (s + 1) * o_peer_rows) of O go
  // through tensor map o_peer_maps[s] (in global memory; it describes destination rank s's output rows for this
  // rank's heads, usually in s's symmetric memory over NVLink) instead of tmO - the epilogue is the way back of the
  // all-to-all.  Null = everything through tmO.
  const CUtensorMap* o_peer_maps;
  void* const* o_peer_ptrs;                      // the destinations' base addresses (device array, behind the maps)
  long long o_peer_stride_h, o_peer_stride_n;    // elements
  int o_peer_rows, o_peer_count;
};

// Token offsets and lengths of the sequence a work item belongs to (the whole tensor when not packed).
struct SeqInfo { int q_off, q_len, k_off, k_len; };
template <bool kVarlen>
__device__ __forceinline__ SeqInfo seq_info(const AttnKernelParams& p, int b) {
  SeqInfo si;
  if constexpr (kVarlen) {
    si.q_off = __ldg(p.cu_q + b); si.q_len = __ldg(p.cu_q + b + 1) - si.q_off;
    si.k_off = __ldg(p.cu_k + b); si.k_len = __ldg(p.cu_k + b + 1) - si.k_off;
  } else {
    si.q_off = 0; si.q_len = p.Nq; si.k_off = 0; si.k_len = p.Nk;
  }
  return si;
}


// ---------------------------------------------------------------------------------------------
// Persistent work list with a dynamic scheduler.  A work item is one pair of Q tiles (256 query rows)
// of one (batch, head).  The TMA producer thread owns the list: a CTA's first item is blockIdx.x,
// every further one comes from an atomic counter (fetched one item ahead, so its latency is hidden),
// is decoded there (the integer divisions stay off the softmax warps) and published to the other
// warps through a small shared-memory ring guarded by mbarriers.
//   non-causal: w -> (bh = w / n_qpairs, q_pair = w % n_qpairs): neighbours in time share K/V in L2
//   causal:     heads in groups of kHeadGroup; inside a group the heavy (late) q pairs come first, so
//               the list is longest-processing-time-first where it matters (the tail of the launch)
//               while only a few heads' K/V are live in L2 at a time.
constexpr int kSchedSlots = 3;
constexpr int kHeadGroup = 8;
// Ticket counters (one pool per translation unit and device, zero-initialised device memory).  Every launch
// zeroes ITS counter with a cudaMemsetAsync on the launch stream just before the kernel (a memset node when
// the launch is being captured), so a launch never depends on what an earlier one left behind.  Which counter:
//   * eager launches: one slot per (stream, host thread for cudaStreamPerThread) - launches on one stream
//     are serialised by the stream, launches on different streams never share a slot;
//   * launches under stream capture: a fresh slot per captured launch that is never handed out again, so a
//     replayed graph cannot collide with eager launches or with another graph (CUDA serialises replays of
//     one executable graph).  The capture slots are finite (kSchedCaptureSlots per process and translation
//     unit); running out is reported as an error, not wrapped around.
constexpr int kSchedStreamSlots = 4096;
constexpr int kSchedCaptureSlots = 61440;
constexpr int kSchedCounters = kSchedStreamSlots + kSchedCaptureSlots;
static __device__ unsigned int g_sched_counter[kSchedCounters];

struct __align__(16) WorkItem {
  int q_pair, h, b, hkv;  // q_pair < 0: end of the list
};
template <bool kVarlen = false>
__device__ __forceinline__ WorkItem decode_work(const AttnKernelParams& p, unsigned int w) {
  WorkItem o;
  if constexpr (kVarlen) {
    // head-major list over the pack's q pairs: w -> (h, global pair) -> (sequence, pair); the number of items is
    // only known on the device
    const int total = __ldg(p.cu_pairs + p.nseq);
    if (total <= 0 || w >= (unsigned int)total * (unsigned int)p.Hq) {
      o.q_pair = -1; o.h = 0; o.b = 0; o.hkv = 0;
      return o;
    }
    const int h = (int)(w / (unsigned int)total);
    const int pg = (int)w - h * total;
    int lo = 0, hi = p.nseq;            // largest s with cu_pairs[s] <= pg
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(p.cu_pairs + mid) <= pg) lo = mid; else hi = mid;
    }
    o.b = lo;
    o.q_pair = pg - __ldg(p.cu_pairs + lo);
    o.h = h;
    o.hkv = h / (p.Hq / p.Hkv);
    return o;
  }
  if (w >= (unsigned int)p.n_work) {
    o.q_pair = -1; o.h = 0; o.b = 0; o.hkv = 0;
    return o;
  }
  int bh, qp;
  if (!p.causal) {
    bh = (int)w / p.n_qpairs;
    qp = (int)w - bh * p.n_qpairs;
  } else {
    const int per_group = kHeadGroup * p.n_qpairs;
    const int g = (int)w / per_group;
    const int r = (int)w - g * per_group;
    const int gsz = min(kHeadGroup, p.B * p.Hq - g * kHeadGroup);
    const int level = r / gsz;
    qp = p.n_qpairs - 1 - level;
    bh = g * kHeadGroup + (r - level * gsz);
  }
  o.q_pair = qp;
  o.b = bh / p.Hq;
  o.h = bh - o.b * p.Hq;
  o.hkv = o.h / (p.Hq / p.Hkv);
  return o;
}
// consumer side of the ring (the slot stays valid until this thread arrives on sched_empty)
__device__ __forceinline__ WorkItem read_work(const WorkItem* ring, uint64_t* sched_full, uint32_t it) {
  mbar_wait(&sched_full[it % kSchedSlots], (it / kSchedSlots) & 1u);
  const int4 v = ld_shared_v4(smem_u32(&ring[it % kSchedSlots]));
  WorkItem o;
  o.q_pair = v.x; o.h = v.y; o.b = v.z; o.hkv = v.w;
  return o;
}
// kv tiles per q tile (causal: stop at the diagonal tile; top-left aligned as the reference,
// /root/reference/csrc/qattn/qk_int_sv_f16_cuda_sm80.cu:218-222)
__device__ __forceinline__ void item_iters(const AttnKernelParams& p, int q_pair, int q_len, int k_len, int (&n_it)[2]) {
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int qs = (q_pair * 2 + t) * kTileQ;
    int kv_end = k_len;
    if (p.causal) kv_end = min(kv_end, qs + kTileQ);
    n_it[t] = (qs < q_len) ? (kv_end + kTileKV - 1) / kTileKV : 0;
  }
}
__device__ __forceinline__ void item_iters(const AttnKernelParams& p, int q_pair, int (&n_it)[2]) {
  item_iters(p, q_pair, p.Nq, p.Nk, n_it);
}
template <bool kVarlen>
__device__ __forceinline__ void item_iters_of(const AttnKernelParams& p, const WorkItem& w, int (&n_it)[2]) {
  if constexpr (kVarlen) {
    const SeqInfo si = seq_info<true>(p, w.b);
    item_iters(p, w.q_pair, si.q_len, si.k_len, n_it);
    // a sequence with queries but no keys still takes ONE kv step, fully masked (lim = -1): P = 0, O = 0, d = 0 and
    // the epilogue writes zeros, as the reference's kernel does (acc = 0, triton/attn_qk_int8_block_varlen.py) - and
    // the pipeline's barriers, which advance per kv step, keep their protocol
    if (si.k_len <= 0) {
#pragma unroll
      for (int t = 0; t < 2; ++t) n_it[t] = ((w.q_pair * 2 + t) * kTileQ < si.q_len) ? 1 : 0;
    }
  } else {
    item_iters(p, w.q_pair, p.Nq, p.Nk, n_it);
  }
}

// Running barrier phases of one softmax warpgroup across work items.
struct WgState {
  uint32_t steps;    // kv steps this warpgroup has finished (phase of s_full / s_free / p_full / p_empty)
  uint32_t items;    // work items it has finished (phase of o_full)
  bool granted;      // warpgroup 1: the first exp turn of the coming item is already granted
};

// What the step loop of one work item needs; prepared ahead of time (under the previous item's last
// P.V) so that the global-load latency of the scales is off the item-to-item critical path.
struct ItemCtx {
  bool valid;
  int n_iter;      // kv steps of this warpgroup's tile (0: tile is beyond Nq)
  int n_pingpong;  // steps both warpgroups execute
  bool next_pingpong;
  int q_row;
  float qs;        // q scale of the row (raw: the load is still in flight when setup returns)
  const float* ksp;
  const uint8_t* mask_row;  // this row's attn_mask (kv column 0), kMask kernels only
  int k_len, ks_stride, ks_count;  // kVarlen kernels: kv length of the sequence, stride / number of its k scales
};
template <int D, bool kVarlen = false>
__device__ __forceinline__ ItemCtx softmax_setup(const AttnKernelParams& p, int t, int row,
                                                 const WorkItem* ring, uint64_t* sched_full, uint32_t it) {
  ItemCtx c;
  const WorkItem wi = read_work(ring, sched_full, it);
  c.valid = wi.q_pair >= 0;
  c.n_iter = c.n_pingpong = 0;
  c.next_pingpong = false;
  c.q_row = 0;
  c.qs = 0.f;
  c.ksp = nullptr;
  c.mask_row = nullptr;
  c.k_len = p.Nk; c.ks_stride = 1; c.ks_count = p.k_scale_per_bh;
  if (!c.valid) return c;
  int n_it[2];
  item_iters_of<kVarlen>(p, wi, n_it);
  c.n_iter = n_it[t];
  c.n_pingpong = min(n_it[0], n_it[1]);
  {  // will both warpgroups alternate in the next item?  (the producer publishes one item ahead)
    const WorkItem nx = read_work(ring, sched_full, it + 1);
    if (nx.q_pair >= 0) {
      int n_nx[2];
      item_iters_of<kVarlen>(p, nx, n_nx);
      c.next_pingpong = min(n_nx[0], n_nx[1]) > 0;
    }
  }
  c.q_row = (wi.q_pair * 2 + t) * kTileQ + row;   // (kVarlen: row inside the sequence)
  if constexpr (kVarlen) {
    // per-block scales of this sequence: [blocks, H], block-major (quant_per_block_varlen.py:46)
    const SeqInfo si = seq_info<true>(p, wi.b);
    const int nqb = (si.q_len + 127) / 128;
    const int tile = min(wi.q_pair * 2 + t, nqb - 1);
    c.qs = __ldg(p.q_scale + (size_t)(__ldg(p.cu_q_scale + wi.b) + tile) * p.Hq + wi.h);
    c.ksp = p.k_scale + (size_t)__ldg(p.cu_k_scale + wi.b) * p.Hkv + wi.hkv;
    c.k_len = si.k_len;
    c.ks_stride = p.Hkv;
    c.ks_count = (si.k_len + 63) / 64;
    return c;
  }
  if (p.mask != nullptr) {
    const long long qr = min(c.q_row, p.Nq - 1);  // rows beyond Nq are computed and clipped: any valid row will do
    c.mask_row = p.mask + (wi.b * p.mask_sb + wi.h * p.mask_sh + qr * p.mask_sm) * (p.mask_kind == 1 ? 1 : 2);
  }
  {
    const int sh = (p.warp_q == 16) ? 4 : 5;
    int idx;
    if (p.gran == 3) idx = (c.q_row >> sh) * 8 + (c.q_row & 7);
    else if (p.gran == 2) idx = c.q_row >> sh;
    else idx = c.q_row >> 7;
    idx = min(idx, p.q_scale_per_bh - 1);
    c.qs = __ldg(p.q_scale + (size_t)(wi.b * p.Hq + wi.h) * p.q_scale_per_bh + idx);  // first use: step 0
  }
  c.ksp = p.k_scale + (size_t)(wi.b * p.Hkv + wi.hkv) * p.k_scale_per_bh;
  prefetch_l1(c.ksp);
  return c;
}


// O row (f32, registers) -> 16-bit row in the 128B-swizzled staging tile.  Packed f32x2 arithmetic:
// y = o * (1/d) [* v_scale] [+ v_mean], two columns per instruction, one F2FP per pair.
template <int D, bool kBf16, bool kScale, bool kMean>
__device__ __forceinline__ void epilogue_convert(const uint32_t (&o)[D], float inv_d, uint32_t vs_addr,
                                                 const float* vmp, uint32_t orow, uint32_t ox) {
  const uint64_t inv2 = pack2(inv_d, inv_d);
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 8) {
    uint32_t packed[4];
#pragma unroll
    for (int i = 0; i < 8; i += 4) {
      uint64_t y0 = fmul2(pack2(__uint_as_float(o[c0 + i]), __uint_as_float(o[c0 + i + 1])), inv2);
      uint64_t y1 = fmul2(pack2(__uint_as_float(o[c0 + i + 2]), __uint_as_float(o[c0 + i + 3])), inv2);
      uint64_t s0 = 0, s1 = 0, m0 = 0, m1 = 0;
      if constexpr (kScale) {
        const int4 f = ld_shared_v4(vs_addr + (uint32_t)(c0 + i) * 4u);
        s0 = pack2(__int_as_float(f.x), __int_as_float(f.y));
        s1 = pack2(__int_as_float(f.z), __int_as_float(f.w));
      }
      if constexpr (kMean) {
        const float4 f = __ldg(reinterpret_cast<const float4*>(vmp + c0 + i));
        m0 = pack2(f.x, f.y);
        m1 = pack2(f.z, f.w);
      }
      if constexpr (kScale && kMean) {
        y0 = ffma2(y0, s0, m0);
        y1 = ffma2(y1, s1, m1);
      } else if constexpr (kScale) {
        y0 = fmul2(y0, s0);
        y1 = fmul2(y1, s1);
      } else if constexpr (kMean) {
        y0 = fadd2(y0, m0);
        y1 = fadd2(y1, m1);
      }
      float a0, a1, a2, a3;
      unpack2(y0, a0, a1);
      unpack2(y1, a2, a3);
      if constexpr (kBf16) {
        __nv_bfloat162 v0 = __floats2bfloat162_rn(a0, a1), v1 = __floats2bfloat162_rn(a2, a3);
        packed[i >> 1] = *reinterpret_cast<uint32_t*>(&v0);
        packed[(i >> 1) + 1] = *reinterpret_cast<uint32_t*>(&v1);
      } else {
        __half2 v0 = __floats2half2_rn(a0, a1), v1 = __floats2half2_rn(a2, a3);
        packed[i >> 1] = *reinterpret_cast<uint32_t*>(&v0);
        packed[(i >> 1) + 1] = *reinterpret_cast<uint32_t*>(&v1);
      }
    }
    // 8 columns = one 16-byte chunk of 64-column block (c0 / 64); chunk c sits at c ^ (row & 7)
    const uint32_t chunk = (uint32_t)((c0 & 63) >> 3);
    const uint32_t addr = orow + (uint32_t)(c0 >> 6) * (kTileQ * 128) + ((chunk ^ ox) << 4);
    st_shared_v4(addr, packed[0], packed[1], packed[2], packed[3]);
  }
}

// Epilogue of one work item: O_t / d [* v_scale] [+ v_mean] -> 16-bit -> shared memory -> bulk tensor
// store; LSE.  The issuer is already running the next item's Q.K^T under it; its first P.V (which
// overwrites O_t) waits for this warpgroup's next P tile, i.e. for the loads done here.
template <int D, bool kFp8, bool kVarlen = false>
__device__ __forceinline__ void softmax_epilogue(const AttnKernelParams& p, int t, int row,
                                                 const WorkItem* ring, uint32_t it, uint32_t tmem_base,
                                                 uint8_t* o_smem, float* vs_smem, const CUtensorMap* tmO,
                                                 uint64_t* o_full, WgState& st, float m_run, float d_run) {
  const uint32_t tO = tmem_base + ((uint32_t)((row >> 5) * 32) << 16) + (uint32_t)(256 + t * D);
  // the item's coordinates come back from the ring slot (still owned by this thread) instead of
  // living in registers through the step loop
  const int4 wv = ld_shared_v4(smem_u32(&ring[it % kSchedSlots]));
  const int h = wv.y, b = wv.z, hkv = wv.w;
  const int q_tile_start = (wv.x * 2 + t) * kTileQ;
  const int q_row = q_tile_start + row;
  // kVarlen: the tile sits at token q_off + q_tile_start of the pack and may end inside the next sequence
  int q_off = 0, rows_valid = kTileQ;
  if constexpr (kVarlen) {
    const SeqInfo si = seq_info<true>(p, b);
    q_off = si.q_off;
    rows_valid = si.q_len - q_tile_start;
  }
#ifdef SAGE_TRACE
  const bool trace_on = blockIdx.x == 0 && row == 0;
#endif
  const float* vsp = p.v_scale ? p.v_scale + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  const float* vmp = p.v_mean ? p.v_mean + (size_t)(b * p.Hkv + hkv) * D : nullptr;
  // per-channel factors: v_scale sits in the ring slot (bulk-copied by the producer together with the
  // entry); the rarer v_mean is only prefetched here and read chunk by chunk below
  if (vmp && (row & 31) < D / 32) prefetch_l1(vmp + (row & 31) * 32);
  SAGE_STAMP(t, st.steps - 1, 10);
  mbar_wait(&o_full[t], st.items & 1u);
  ++st.items;
  tc_fence_after();
  SAGE_STAMP(t, st.steps - 1, 8);
  const float inv_d = (d_run > 0.f) ? rcp_approx(d_run) : 0.f;  // d = 0: every key of the row masked (attn_mask) -> zeros
  // The tile goes out through shared memory and ONE bulk tensor store per 64-column block (a thread
  // owns a row, so direct stores would be 16 bytes per thread at a row pitch: measured 5-7 k cycles
  // per item).  The previous item's store must have finished reading the staging tile first.
  if (row == 0) tma_store_wait_read();
  named_bar_sync(3 + t, kTileQ);
  SAGE_STAMP(t, st.steps - 1, 11);
  const uint32_t orow = smem_u32(o_smem) + (uint32_t)row * 128u;
  const uint32_t ox = (uint32_t)(row & 7);
  // all of O_t in one go: the four loads overlap, one wait
  uint32_t oall[D];
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 32)
    tmem_ld32(tO + c0, *reinterpret_cast<uint32_t(*)[32]>(&oall[c0]));
  tc_wait_ld();
  // one specialised conversion loop per (output type, v_scale?, v_mean?) instead of per-element selects
  const uint32_t vs_addr = smem_u32(vs_smem);
  const int variant = (p.out_bf16 ? 4 : 0) | (vsp ? 2 : 0) | (vmp ? 1 : 0);
  switch (variant) {
    case 0: epilogue_convert<D, false, false, false>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 1: epilogue_convert<D, false, false, true>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 2: epilogue_convert<D, false, true, false>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 3: epilogue_convert<D, false, true, true>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 4: epilogue_convert<D, true, false, false>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 5: epilogue_convert<D, true, false, true>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    case 6: epilogue_convert<D, true, true, false>(oall, inv_d, vs_addr, vmp, orow, ox); break;
    default: epilogue_convert<D, true, true, true>(oall, inv_d, vs_addr, vmp, orow, ox); break;
  }
  // sequence-parallel output scatter: a tile whose rows belong to TWO destination ranks goes out row by row (a bulk
  // tensor store cannot start at a negative row of the second destination's map: illegal instruction, measured)
  bool straddle = false;
  int peer0 = 0;
  if constexpr (!kVarlen) {
    if (p.o_peer_maps != nullptr) {
      peer0 = q_tile_start / p.o_peer_rows;
      straddle = (min(q_tile_start + kTileQ, p.Nq) - 1) / p.o_peer_rows != peer0;
      if (straddle && q_row < p.Nq) {
        const int s_ = q_row / p.o_peer_rows;
        uint8_t* dst = reinterpret_cast<uint8_t*>(p.o_peer_ptrs[s_]) +
                       ((long long)(q_row - s_ * p.o_peer_rows) * p.o_peer_stride_n + (long long)h * p.o_peer_stride_h) * 2;
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 8) {
          const uint32_t chunk = (uint32_t)((c0 & 63) >> 3);
          const int4 v = ld_shared_v4(orow + (uint32_t)(c0 >> 6) * (kTileQ * 128) + ((chunk ^ ox) << 4));
          *reinterpret_cast<int4*>(dst + c0 * 2) = v;
        }
      }
    }
  }
  if constexpr (kVarlen) {
    if (rows_valid < kTileQ) {
      // last tile of a sequence: a bulk tensor store would run into the next sequence's rows, so every thread
      // copies its own row (which it has just written to the staging tile) to global memory
      if (row < rows_valid) {
        uint8_t* dst = reinterpret_cast<uint8_t*>(p.o) +
                       ((long long)(q_off + q_row) * p.o_stride_n + (long long)h * p.o_stride_h) * 2;
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 8) {
          const uint32_t chunk = (uint32_t)((c0 & 63) >> 3);
          const int4 v = ld_shared_v4(orow + (uint32_t)(c0 >> 6) * (kTileQ * 128) + ((chunk ^ ox) << 4));
          *reinterpret_cast<int4*>(dst + c0 * 2) = v;
        }
      }
    }
  }
  fence_proxy_async_smem();
  named_bar_sync(3 + t, kTileQ);
  if (row == 0 && (!kVarlen || rows_valid >= kTileQ) && !straddle) {
    if (!kVarlen && p.o_peer_maps != nullptr) {
      // all of the tile's rows (< Nq) belong to destination peer0: its map clips the rows beyond its o_peer_rows
      if (peer0 < p.o_peer_count) {
#pragma unroll
        for (int c = 0; c < D / 64; ++c)
          tma_store_4d(p.o_peer_maps + peer0, o_smem + c * (kTileQ * 128), c * 64, q_tile_start - peer0 * p.o_peer_rows, h, 0);
      }
    } else {
#pragma unroll
      for (int c = 0; c < D / 64; ++c)
        tma_store_4d(tmO, o_smem + c * (kTileQ * 128), c * 64, q_off + q_tile_start, h, kVarlen ? 0 : b);
    }
    tma_store_commit();
  }
  const bool store = q_row < p.Nq;
  tc_fence_before();
  SAGE_STAMP(t, st.steps - 1, 9);
  if (p.lse && store)
    p.lse[(size_t)(b * p.Hq + h) * p.Nq + q_row] =
        lg2_approx(d_run) + m_run - (kFp8 ? kFp8Offset : 0.f);
}


// ==========================================================================================
// Host side: TMA descriptors, per-device state, argument checks
// ==========================================================================================
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

// SMs of the current device (the persistent grid runs one CTA on each)
static int sm_count() {
  static std::atomic<int> cached[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// cudaFuncSetAttribute(max dynamic shared memory) once per (kernel instantiation = call site, device); racing
// threads may both set it, which is harmless.
struct SmemAttrOnce {
  std::atomic<bool> done[64];
  template <typename K>
  cudaError_t ensure(K kern, int bytes) {
    int dev = 0;
    cudaGetDevice(&dev);
    const bool tracked = dev >= 0 && dev < 64;
    if (tracked && done[dev].load(std::memory_order_acquire)) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && tracked) done[dev].store(true, std::memory_order_release);
    return e;
  }
};

// The scheduler counter of one launch on `stream` (see g_sched_counter); nullptr + message on failure.
static unsigned int* acquire_sched_counter(cudaStream_t stream) {
  struct Key {
    int dev; cudaStream_t stream; std::thread::id tid;
    bool operator<(const Key& o) const {
      if (dev != o.dev) return dev < o.dev;
      if (stream != o.stream) return stream < o.stream;
      return tid < o.tid;
    }
  };
  static std::mutex mu;
  static std::map<Key, int> by_stream;
  static int next_stream_slot = 0;
  static int next_capture_slot[64] = {0};
  static unsigned int* base[64] = {nullptr};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) {
    cudaGetLastError();
    cap = cudaStreamCaptureStatusNone;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (base[dev] == nullptr) {
    void* ptr = nullptr;
    if (cudaGetSymbolAddress(&ptr, g_sched_counter) != cudaSuccess) {
      fail_arg("cudaGetSymbolAddress(g_sched_counter) failed");
      return nullptr;
    }
    base[dev] = static_cast<unsigned int*>(ptr);
  }
  int slot;
  if (cap == cudaStreamCaptureStatusActive) {
    if (next_capture_slot[dev] >= kSchedCaptureSlots) {
      fail_arg("more than %d attention launches captured into CUDA graphs on device %d", kSchedCaptureSlots, dev);
      return nullptr;
    }
    slot = kSchedStreamSlots + next_capture_slot[dev]++;
  } else {
    Key key{dev, stream, stream == cudaStreamPerThread ? std::this_thread::get_id() : std::thread::id()};
    auto it = by_stream.find(key);
    if (it == by_stream.end()) {
      // more than kSchedStreamSlots distinct live streams would share slots; handles of destroyed streams are
      // recycled by the driver, so in practice the map stays small
      it = by_stream.emplace(key, next_stream_slot++ % kSchedStreamSlots).first;
    }
    slot = it->second;
  }
  return base[dev] + slot;
}


static int check_attn_args(const sage_attn_args* a) {
  SAGE_REQUIRE(a != nullptr, "null args");
  SAGE_REQUIRE(a->D == 64 || a->D == 128, "head_dim must be 64 or 128 after padding (got %d)", a->D);
  SAGE_REQUIRE(a->B > 0 && a->Hq > 0 && a->Hkv > 0 && a->Nq > 0 && a->Nk > 0, "empty problem");
  SAGE_REQUIRE(a->Hq % a->Hkv == 0, "num_qo_heads (%d) must be divisible by num_kv_heads (%d)", a->Hq, a->Hkv);
  SAGE_REQUIRE(a->gran >= 1 && a->gran <= 3, "unsupported qk_quant_gran %d", a->gran);
  SAGE_REQUIRE(a->gran == 1 || a->warp_q == 16 || a->warp_q == 32, "warp_q must be 16 or 32");
  SAGE_REQUIRE(a->q8.ptr && a->k8.ptr && a->v.ptr && a->o.ptr && a->q_scale && a->k_scale, "null tensor pointer");
  SAGE_REQUIRE((reinterpret_cast<uintptr_t>(a->k_scale) & 15) == 0, "k_scale must be 16-byte aligned (it is read with 16-byte loads)");
  SAGE_REQUIRE(a->o.stride_n % 8 == 0 && a->o.stride_h % 8 == 0 && a->o.stride_b % 8 == 0 &&
               (reinterpret_cast<uintptr_t>(a->o.ptr) & 15) == 0, "output rows must be 16-byte aligned");
  SAGE_REQUIRE(a->B <= 65535 && a->Hq <= 65535, "batch / heads exceed grid limits");
  if (a->cu_seqlens_q != nullptr) {
    SAGE_REQUIRE(a->cu_seqlens_k && a->cu_seqlens_q_scale && a->cu_seqlens_k_scale && a->cu_pairs && a->nseq > 0,
                 "varlen: all five prefix arrays and nseq are needed");
    SAGE_REQUIRE(a->B == 1 && a->gran == SAGE_GRAN_PER_BLOCK, "varlen: packed tensors are passed as B = 1 with per-block scales");
    SAGE_REQUIRE(a->mask == nullptr && a->lse == nullptr && a->v_scale == nullptr && a->v_mean == nullptr && a->d_og == 0,
                 "varlen: attn_mask, LSE, fp8 V, v_mean and head_dim folding are not supported");
  }
  if (a->mask != nullptr) {
    SAGE_REQUIRE(a->mask_kind >= 1 && a->mask_kind <= 3, "attn_mask kind must be 1 (bool), 2 (f16) or 3 (bf16), got %d", a->mask_kind);
    SAGE_REQUIRE(!a->is_causal, "attn_mask must be null for causal attention");
    const long long esz = a->mask_kind == 1 ? 1 : 2;
    SAGE_REQUIRE((reinterpret_cast<uintptr_t>(a->mask) & 15) == 0 && (a->mask_stride_b * esz) % 16 == 0 &&
                 (a->mask_stride_h * esz) % 16 == 0 && (a->mask_stride_m * esz) % 16 == 0,
                 "attn_mask rows must be 16-byte aligned (pad the last dimension to a multiple of 128)");
    SAGE_REQUIRE(a->mask_stride_m >= (a->Nk + 127) / 128 * 128,
                 "attn_mask rows must hold ceil128(kv_len) = %d elements (got row stride %lld)", (a->Nk + 127) / 128 * 128,
                 (long long)a->mask_stride_m);
  }
  return 0;
}


}  // namespace sage
