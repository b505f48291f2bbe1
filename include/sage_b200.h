/* sage_b200.h - C ABI of libsage_b200.so (sm_100a SageAttention hot path).
 *
 * Drop-in boundary for the reference's pybind extensions.  Every entry point takes plain device
 * pointers, sizes, element strides and a CUDA stream; nothing is allocated by the library and no
 * torch type crosses the boundary.  Return value: 0 = ok, <0 = invalid argument (message in
 * sage_last_error()), >0 = cudaError_t of the failing runtime call / launch.
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   sage_quant_int8              <- _fused.quant_per_block_int8_cuda (2 overloads),
 *                                   quant_per_block_int8_fuse_sub_mean_cuda, quant_per_warp_int8_cuda
 *                                   (csrc/fused/pybind.cpp:23-26, csrc/fused/fused.cu:429-768)
 *   sage_quant_per_thread_int8   <- triton per_thread_int8 (sageattention/triton/quant_per_thread.py:158-208)
 *   sage_sub_mean                <- _fused.sub_mean_cuda (csrc/fused/fused.cu:770-848)
 *   sage_per_channel_fp8         <- _fused.transpose_pad_permute_cuda + scale_fuse_quant_cuda /
 *                                   mean_scale_fuse_quant_cuda (csrc/fused/fused.cu:850-1083)
 *   sage_seq_mean                <- torch `k.mean(dim=seq)` / `v.mean(dim=seq)` (sageattention/core.py:612, quant.py:215)
 *   sage_attn_i8_f16             <- _qattn_sm80.qk_int8_sv_f16_accum_{f32,f16,f16_inst_buf}_attn,
 *                                   ..._fuse_v_mean_attn (csrc/qattn/pybind_sm80.cpp, qk_int_sv_f16_cuda_sm80.cu:674-1379)
 *   sage_attn_i8_f8              <- _qattn_sm89.qk_int8_sv_f8_accum_* (7 variants, csrc/qattn/pybind_sm89.cpp)
 *   sage_prepass_qk_int8         <- the whole Q/K prologue of the operator with smooth_k (sageattention/core.py:611-632,
 *                                   794-812: k.mean, per_warp_int8 / per_thread_int8 for q and k) in one launch
 *   sage_prepass_qkv_fp8         <- the whole prologue of the fp8-P.V operator without smooth_v (sageattention/core.py:
 *                                   794-850: k.mean, per_warp_int8 / per_thread_int8, V padding + per_channel_fp8) in one launch
 *   sage_smooth_quant_k_int8     <- `km = k.mean(dim=seq)` followed by the K half of per_warp_int8 /
 *                                   per_thread_int8 (sageattention/core.py:611-632), fused in one launch
 */
#ifndef SAGE_B200_H_
#define SAGE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes */
#define SAGE_DTYPE_F16 0
#define SAGE_DTYPE_BF16 1
/* QuantGranularity codes, identical to the reference's enum (csrc/qattn/attn_utils.cuh:44-50) */
#define SAGE_GRAN_PER_BLOCK 1
#define SAGE_GRAN_PER_WARP 2
#define SAGE_GRAN_PER_THREAD 3

/* A 4-D activation tensor [B, H, N, D] addressed through element strides, so both of the
 * reference's layouts ("HND" and "NHD", core.py:585) and transposed views are one code path.
 * stride_d is always 1 (the reference asserts stride(-1) == 1, core.py:604). */
typedef struct sage_tensor4 {
  void* ptr;
  int64_t stride_b, stride_h, stride_n; /* in elements */
} sage_tensor4;

const char* sage_last_error(void);
int sage_version(void);
/* sha256 (hex) of the csrc/ and include/ sources the library was built from, "unknown" for a hand build */
const char* sage_source_hash(void);

/* km[b,h,d] = dtype( (sum_n x[b,h,n,d]) * (1/N) ); out is [B,H,D] contiguous, same dtype as x.
 * Deterministic two-level reduction; workspace of sage_seq_mean_workspace_bytes() bytes. */
size_t sage_seq_mean_workspace_bytes(int B, int H, int N, int D);
int sage_seq_mean(sage_tensor4 x, void* out, int B, int H, int N, int D, int dtype, void* workspace,
                  void* stream);

/* Packed variable-length batches (sageattn_varlen): per-block INT8 quantisation of q [Tq,Hq,D] (pre-scaled by
 * sm_scale_q = sm_scale * log2 e) and k [Tk,Hkv,D] (minus km [Hkv,D] when km != NULL) in ONE launch, bit-identical to
 * the reference's Triton kernel (/root/reference/sageattention/triton/quant_per_block_varlen.py:20-107): blocks of 128
 * (q) / 64 (k) rows restarting at every sequence, scale = amax / 127 with no floor, round half away from zero.
 * q_scale / k_scale: [blocks, H] like the reference's, sequence b starting at row cu_seqlens_*_scale[b]; those prefix
 * arrays [nseq+1] and cu_pairs [nseq+1] (prefix of ceil(q_len/256), the attention kernel's work list) are OUTPUTS of
 * this launch.  The caller sizes the scale buffers for the upper bound ceil(T/BLK) + nseq rows: no length is needed on
 * the host.  Strides in elements; outputs contiguous. */
int sage_quant_per_block_varlen_int8(const void* q, const void* k, const void* km, int8_t* q_out, int8_t* k_out,
                                     float* q_scale, float* k_scale, const int* cu_seqlens_q, const int* cu_seqlens_k,
                                     int* cu_seqlens_q_scale, int* cu_seqlens_k_scale, int* cu_pairs, int nseq, int Hq,
                                     int Hkv, int D, int dtype, int max_seqlen_q, int max_seqlen_k, int64_t q_stride_n,
                                     int64_t q_stride_h, int64_t k_stride_n, int64_t k_stride_h, float sm_scale_q,
                                     void* stream);

/* sage_seq_mean without a workspace argument and without any allocation.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic and
 * bit-identical result as sage_seq_mean; the reduction scratch is a static device array of
 * SAGE_SEQ_MEAN_STATIC_MAX_BYTES inside the library, which every GPU that loads the library holds until exit.
 * The call makes no cudaMalloc / cudaFree / allocating library call; the caller still passes x and out.
 * Limit: B*H*ceil(N/256)*3*D*4 + B*H*4 + 256 bytes must fit the array, else -1 with a message naming both sizes
 * (Wan2.1 K, B1 H40 N75600 D128, needs 18.2 MB; HunyuanVideo, B1 H24 N119056 D128, 17.2 MB).
 * All calls on a device SHARE the array: calls that overlap on the GPU (different streams, concurrent graph
 * replays) race; serialise them on one stream or use sage_seq_mean with a workspace each.  The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress): make it once at set-up, outside capture. */
#define SAGE_SEQ_MEAN_STATIC_MAX_BYTES (32u << 20)
int sage_seq_mean_static(sage_tensor4 x, void* out, int B, int H, int N, int D, int dtype, void* stream);

/* INT8 quantisation of rows in groups of `block` tokens (block in {16,32,64,128}).
 *   x' = float(x) [- float(mean[b,h,d])] [* sm_scale]   (mean may be NULL; has_sm_scale 0/1)
 *   amax = max(1e-7, max|x'|); scale = amax/127; code = cvt.rni.sat.s8(x' * (127/amax))
 * out: int8 tensor (strides given), scale: [B,H,scale_per_bh] contiguous fp32 with
 * scale_per_bh >= ceil(N/block) (per_warp_int8 allocates ceil(N/BLKQ)*(BLKQ/WARPQ), quant.py:170). */
int sage_quant_int8(sage_tensor4 x, const void* mean, sage_tensor4 out, float* scale, int B, int H,
                    int N, int D, int dtype, int block, int scale_per_bh, int has_sm_scale,
                    float sm_scale, void* stream);

/* Per-thread granularity (the reference's default, Triton arithmetic):
 *   scale = max|x|/127 + 1e-7 ; code = trunc(x/scale + 0.5*sign)
 *   is_key = 0: groups {r, r+8, ...} of each `tile`-row block (tile = WARPQ: 16 or 32), 8 scales per tile
 *   is_key = 1: groups {8i+2c, 8i+2c+1} of each 64-row block, 4 scales per block; x is first
 *               replaced by dtype(x - mean) when mean != NULL.
 * scale: [B,H,ceil(N/128)*(128/tile)*8] (query) or [B,H,ceil(N/64)*4] (key). */
int sage_quant_per_thread_int8(sage_tensor4 x, const void* mean, sage_tensor4 out, float* scale,
                               int B, int H, int N, int D, int dtype, int is_key, int tile,
                               void* stream);

/* sage_quant_per_thread_int8 with an explicit distance scale_per_bh (>= the entries N rows produce) between the scale
 * rows of consecutive (b, h): a caller that quantises one sequence in several calls over row ranges that start on 128-row
 * boundaries points each call at its columns of one [B, H, scale_per_bh] array (same reference arithmetic,
 * /root/reference/sageattention/triton/quant_per_thread.py:22-102). */
int sage_quant_per_thread_int8_strided(sage_tensor4 x, const void* mean, sage_tensor4 out, float* scale,
                                       int B, int H, int N, int D, int dtype, int is_key, int tile,
                                       int scale_per_bh, void* stream);

/* K smoothing and quantisation in one launch (a thread-block cluster per (b,h) slice reads K once
 * from HBM): mean_out[b,h,:] = dtype(mean over the sequence), then the codes and scales that
 * sage_quant_int8(block = 64, mean) (per_thread = 0) or sage_quant_per_thread_int8(is_key = 1, mean)
 * (per_thread = 1) produce for that mean, bit for bit.
 * scale: [B,H,scale_per_bh], scale_per_bh >= ceil(N/64) (per_thread = 0) or ceil(N/64)*4. */
int sage_smooth_quant_k_int8(sage_tensor4 k, void* mean_out, sage_tensor4 out, float* scale, int B, int H,
                             int N, int D, int dtype, int per_thread, int scale_per_bh, void* stream);

/* The Q/K prologue in one launch: what sage_smooth_quant_k_int8(k) and the Q half
 * (sage_quant_per_thread_int8(is_key = 0, tile = warp_q) if per_thread, else sage_quant_int8(block =
 * warp_q)) produce, bit for bit.  q: [B,Hq,Nq,D], k: [B,Hkv,Nk,D] (strides given); k_mean_out [B,Hkv,D].
 * q_scale: [B,Hq,q_scale_per_bh] (>= ceil(Nq/128)*(128/warp_q)*8 per_thread, >= ceil(Nq/warp_q) else),
 * k_scale: [B,Hkv,k_scale_per_bh] (>= ceil(Nk/64)*4 per_thread, >= ceil(Nk/64) else).
 * d_og (0 = D): real head_dim of q / k when it is not 64 or 128 (a multiple of 8, <= D): the input rows
 * hold d_og elements, the codes and the mean are written D wide with zeros beyond d_og - the caller-side
 * F.pad of core.py:590-601 without its three extra passes over q, k, v. */
int sage_prepass_qk_int8(sage_tensor4 q, sage_tensor4 k, sage_tensor4 q_out, sage_tensor4 k_out,
                         float* q_scale, float* k_scale, void* k_mean_out, int B, int Hq, int Hkv, int Nq,
                         int Nk, int D, int dtype, int per_thread, int warp_q, int q_scale_per_bh,
                         int k_scale_per_bh, int d_og, void* stream);

/* sage_prepass_qk_int8 plus, from the same launch, v16 = half(v): the f16 copy the f16-P.V operator makes of a
 * bf16 V (/root/reference/sageattention/core.py:633 `v.to(torch.float16)`, a separate eager kernel there and in
 * round 1 here).  v: [B,Hkv,Nk,d_og] in the layout of k; v16_out: same shape, f16, rows 16-byte aligned.  Codes,
 * scales and mean are bit-identical to sage_prepass_qk_int8; v16 is bit-identical to the torch cast. */
int sage_prepass_qk_int8_v16(sage_tensor4 q, sage_tensor4 k, sage_tensor4 q_out, sage_tensor4 k_out,
                             float* q_scale, float* k_scale, void* k_mean_out, sage_tensor4 v,
                             sage_tensor4 v16_out, int B, int Hq, int Hkv, int Nq, int Nk, int D, int dtype,
                             int per_thread, int warp_q, int q_scale_per_bh, int k_scale_per_bh, int d_og,
                             void* stream);

/* The Q/K/V prologue of the fp8-P.V operator (no smooth_v) in one launch: what sage_prepass_qk_int8(q, k) and
 * sage_per_channel_fp8(v, smooth_v = 0) produce, bit for bit (same arguments; v is [B,Hkv,Nk,D], n_stat as for
 * sage_per_channel_fp8, no workspace needed). */
int sage_prepass_qkv_fp8(sage_tensor4 q, sage_tensor4 k, sage_tensor4 q_out, sage_tensor4 k_out,
                         float* q_scale, float* k_scale, void* k_mean_out, sage_tensor4 v, uint8_t* v_out,
                         int64_t vo_stride_b, int64_t vo_stride_h, int64_t vo_stride_d, float* v_scale,
                         int B, int Hq, int Hkv, int Nq, int Nk, int n_stat, int D, int dtype,
                         int per_thread, int warp_q, int q_scale_per_bh, int k_scale_per_bh,
                         float scale_max, int d_og, void* stream);

/* out(f16)[b,h,n,d] = half( x - mean ) in the input type's packed arithmetic. */
int sage_sub_mean(sage_tensor4 x, const void* mean, sage_tensor4 out, int B, int H, int N, int D,
                  int dtype, void* stream);

/* V -> e4m3 [B,H,D,Npad] (16-token permutation of the reference), per-channel scale [B,H,D] fp32
 * and (smooth_v) mean [B,H,D] fp32.  v_out strides: (b, h, d) with n contiguous.
 * n_stat >= N is the token count the reference kernel would see: core.py:822-828 zero-pads V to a
 * multiple of 128 before quantising, so Npad = ceil64(n_stat) and the smooth_v mean divides by
 * ceil16(n_stat); rows in [N, n_stat) are treated as zeros without being materialised.
 * workspace: sage_per_channel_fp8_workspace_bytes(). */
size_t sage_per_channel_fp8_workspace_bytes(int B, int H, int N, int D);
int sage_per_channel_fp8(sage_tensor4 v, uint8_t* v_out, int64_t vo_stride_b, int64_t vo_stride_h,
                         int64_t vo_stride_d, float* v_scale, float* v_mean, int B, int H, int N,
                         int n_stat, int D, int dtype, float scale_max, int smooth_v, void* workspace,
                         void* stream);

/* sage_per_channel_fp8 with smooth_v = 0 as two flat launches (streaming channel range, then the conversion tiles) instead
 * of one cluster per (b, h) slice; same outputs.  For long sequences of few heads (sequence-parallel DiT shards) the
 * cluster launch leaves most SMs idle; sage_sageattn_fwd switches to this above 65 536 kv rows.  Same workspace. */
int sage_per_channel_fp8_flat(sage_tensor4 v, uint8_t* v_out, int64_t vo_stride_b, int64_t vo_stride_h,
                              int64_t vo_stride_d, float* v_scale, int B, int H, int N, int n_stat, int D, int dtype,
                              float scale_max, void* workspace, void* stream);

/* Sequence-parallel callers that quantise BEFORE the all-to-all (the reference's only sequence-parallel
 * precedent is the xDiT monkey patch of /root/reference/example/parallel_sageattn_cogvideo.py:43-52, which ships the
 * 16-bit activations): the partial statistics of this rank's sequence shard x [B,H,N,D] - raw f32 column sums
 * (sum_out [B,H,D], for the global mean of K: the caller adds the ranks' sums, multiplies by 1/N_total and rounds to the
 * input dtype, which is sage_seq_mean's arithmetic) and max|x| per channel (absmax_out [B,H,D], for the global range of V:
 * the caller takes the maximum over ranks).  Either output may be NULL.  Workspace: sage_seq_mean_workspace_bytes(). */
int sage_col_sum_absmax(sage_tensor4 x, float* sum_out, float* absmax_out, int B, int H, int N, int D, int dtype,
                        void* workspace, void* stream);
/* The conversion half of sage_per_channel_fp8_flat for a channel range the CALLER provides (absmax [B,H,D] f32 = max|v|
 * over the whole sequence, of which v may be one shard): e4m3 codes in the reference's [B,H,D,Npad] layout with its
 * 16-token permutation (/root/reference/csrc/fused/fused.cu:262-427); v_scale_out (or NULL) receives absmax / scale_max.
 * With the full sequence's range the codes and scales are bit-identical to sage_per_channel_fp8's. */
int sage_v_fp8_from_absmax(sage_tensor4 v, uint8_t* v_out, int64_t vo_stride_b, int64_t vo_stride_h, int64_t vo_stride_d,
                           const float* absmax, float* v_scale_out, int B, int H, int N, int n_stat, int D, int dtype,
                           float scale_max, void* stream);

/* The two calls above with ONE OUTPUT BASE POINTER PER "BATCH" (B <= 16): in the quantised-payload sequence-parallel path
 * the batch dimension is the destination rank, and with out_ptrs / scale_ptrs (HOST arrays of B device pointers) aimed at
 * the peers' symmetric-memory receive buffers the quantiser's stores travel over NVLink - quantiser and all-to-all in
 * one launch, no NCCL kernel on the inbound path.  No counterpart in the reference (its sequence-parallel example ships
 * 16-bit activations through xDiT, example/parallel_sageattn_cogvideo.py:43-52).  Codes of (b, h, n) go to
 * out_ptrs[b] + h*out_stride_h + n*out_stride_n, scales to scale_ptrs[b] + h*scale_per_bh. */
int sage_quant_per_thread_int8_peers(sage_tensor4 x, const void* mean, void* const* out_ptrs, int64_t out_stride_h,
                                     int64_t out_stride_n, float* const* scale_ptrs, int B, int H, int N, int D,
                                     int dtype, int is_key, int tile, int scale_per_bh, void* stream);
int sage_v_fp8_from_absmax_peers(sage_tensor4 v, void* const* out_ptrs, int64_t vo_stride_h, int64_t vo_stride_d,
                                 const float* absmax, float* v_scale_out, int B, int H, int N, int n_stat, int D,
                                 int dtype, float scale_max, void* stream);

/* Attention kernels.  q8/k8: int8 [B,H,N,D] via strides; scales laid out exactly as the reference
 * quantizers produce them; o: f16/bf16 via strides; lse: [B,Hq,Nq] fp32 or NULL (base-2, as the
 * reference kernels return it). */
typedef struct sage_attn_args {
  sage_tensor4 q8, k8, v, o;
  const float* q_scale;
  const float* k_scale;
  const float* v_scale; /* fp8: [B,Hkv,D] or NULL */
  const float* v_mean;  /* [B,Hkv,D] fp32 or NULL */
  float* lse;           /* or NULL */
  int B, Hq, Hkv, Nq, Nk, D;
  int npad_v;      /* fp8: padded kv length of v (multiple of 64); v strides are (b,h,d) rows */
  int gran;        /* SAGE_GRAN_* */
  int warp_q;      /* rows per q-scale group for per_warp / per_thread: 16 or 32 */
  int is_causal;
  int out_dtype;   /* SAGE_DTYPE_* */
  float sm_scale;
  void* debug_s;   /* NULL, or int32 [B,Hq,ceil256(Nq),128]: raw QK^T of kv tile 0 (test hook) */
  int d_og;        /* 0 = D, else the real head_dim (multiple of 8, <= D) of o and of an f16 v: their rows hold
                      d_og elements; TMA zero-fills / clips the columns up to D (no F.pad, no output slice) */
  int max_ctas;    /* 0 = one CTA per SM; else an upper bound on the persistent grid, so that a concurrent kernel
                      on another stream (an NCCL collective under sequence parallelism) finds free SMs */
  const void* mask;  /* NULL, or attn_mask [B|1, H|1, Nq, Lpad] (Lpad >= ceil128(Nk), the padding is never used; broadcast
                        dims have stride 0; rows 16-byte aligned).  Reference: sageattn_qk_int8_pv_fp16_triton's attn_mask,
                        /root/reference/sageattention/core.py:249-251, triton/attn_qk_int8_per_thread.py:39-72.  Not with is_causal. */
  int mask_kind;     /* 1: bool as uint8 (non-zero = attend); 2 / 3: f16 / bf16, ADDED to the base-2 logits as it is (the
                        reference does not scale it by log2 e) */
  int64_t mask_stride_b, mask_stride_h, mask_stride_m; /* in elements */
  /* Packed variable-length batches (sageattn_varlen, /root/reference/sageattention/core.py:363-477; f16 P.V only).
     NULL cu_seqlens_q = fixed-length call.  Otherwise q8 / k8 / v / o are [tokens, H, D] passed as B = 1, Nq / Nk = the
     token counts, gran = SAGE_GRAN_PER_BLOCK; the five int32 device arrays [nseq + 1] are the caller's cu_seqlens and
     the outputs of sage_quant_per_block_varlen_int8; q_scale / k_scale are its [blocks, H] scale arrays.  ONE launch
     covers the pack: the persistent work list is derived from cu_pairs on the device (no length is read by the host). */
  const int* cu_seqlens_q; const int* cu_seqlens_k; const int* cu_seqlens_q_scale; const int* cu_seqlens_k_scale;
  const int* cu_pairs;
  int nseq;
  /* Sequence-parallel output scatter (no counterpart in the reference; B = 1, fixed length).  NULL o_peer_ptrs = all of O
     goes to `o`.  Otherwise rows [s * o_peer_rows, (s + 1) * o_peer_rows) of O are stored to o_peer_ptrs[s] (HOST array of
     o_peer_count <= 16 device pointers, normally the destination ranks' symmetric-memory output buffers, reached over
     NVLink) as [o_peer_rows, Hq, D] with the given element strides - the kernel's TMA-store epilogue is the way back of
     the Ulysses all-to-all (a tile whose rows belong to two destinations is stored row by row).  o_peer_maps: device
     scratch of SAGE_O_PEER_SCRATCH_BYTES, 64-byte aligned, which the call fills (on the stream) with the destinations'
     tensor maps and addresses; it must stay untouched until the kernel has finished; `o` must still be a valid tensor. */
#define SAGE_O_PEER_SCRATCH_BYTES (16 * 128 + 16 * 8)
  void* const* o_peer_ptrs;
  int o_peer_count, o_peer_rows;
  int64_t o_peer_stride_h, o_peer_stride_n;
  void* o_peer_maps;
} sage_attn_args;

int sage_attn_i8_f16(const sage_attn_args* a, void* stream);
int sage_attn_i8_f8(const sage_attn_args* a, void* stream);

/* The whole operator behind ONE call and ONE workspace: quantisation prologue (smooth_k on, smooth_v off) + attention,
 * for the default paths of sageattn / sageattn_qk_int8_pv_fp8_cuda (/root/reference/sageattention/core.py:656-905) and
 * sageattn_qk_int8_pv_fp16_cuda (core.py:480-653).  It launches exactly what sage_prepass_qkv_fp8 /
 * sage_prepass_qk_int8[_v16] followed by sage_attn_i8_f8 / sage_attn_i8_f16 launch (results bit-identical); what it
 * saves is host time: the caller makes two allocations (workspace, o) and one foreign call instead of 6-8 and 2-3.
 * q, k, v, o: the caller's tensors ([B,H,N,d_og] or [B,N,H,d_og], any strides with contiguous rows).  workspace:
 * sage_sageattn_fwd_workspace_bytes() bytes, 256-byte aligned, contents undefined afterwards.  km_out: NULL, or
 * [B,Hkv,D] in the input dtype to receive mean(K) (the LSE correction needs it, core.py:613-617).  lse: NULL or f32
 * [B,Hq,Nq] (base 2, uncorrected).  ev_start / ev_stop: NULL, or cudaEvent_t recorded on the stream around the attention
 * launch alone (bench.py's kernel-only figure). */
typedef struct {
  sage_tensor4 q, k, v, o;
  void* workspace;
  size_t workspace_bytes;
  void* km_out;
  float* lse;
  void* ev_start;
  void* ev_stop;
  int B, Hq, Hkv, Nq, Nk;
  int D;          /* 64 or 128: head_dim after padding */
  int d_og;       /* 0 = D, else the real head_dim (multiple of 8) */
  int dtype;      /* SAGE_DTYPE_* of q, k, v, o */
  int layout_hnd; /* 1: [B,H,N,D], 0: [B,N,H,D] (the reference's _tensor_layout) */
  int pv_fp8;     /* 1: e4m3 P.V (sageattn), 0: f16 P.V */
  int per_thread; /* 1: per_thread, 0: per_warp */
  int warp_q;     /* 32, or 16 */
  int is_causal;
  int max_ctas;   /* 0 = one CTA per SM */
  float sm_scale;
  float v_scale_max; /* fp8 P.V: 448 or 2.25 (core.py:818-820) */
} sage_fwd_args;
size_t sage_sageattn_fwd_workspace_bytes(const sage_fwd_args* a);
int sage_sageattn_fwd(const sage_fwd_args* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SAGE_B200_H_ */
