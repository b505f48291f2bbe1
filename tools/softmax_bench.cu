// Stand-alone throughput test of the softmax row code (no MMA, no pipeline barriers): every
// softmax thread repeatedly loads 128 s32 logits of its row from TMEM, runs the production
// softmax_row<> (max, dequant, exp2, sum, pack) and stores P back to TMEM.  Reports SM cycles per
// "step pair" (2 Q tiles x 128 x 128 logits), the unit the attention kernel's tensor work
// (2 x (QK 256 + PV 256) = 1024 cycles at hd128 fp8) has to be compared with.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iinclude -Isageattention_b200/csrc \
//        -o tools/_bin/softmax_bench tools/softmax_bench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include "softmax_sm100.cuh"

using namespace sage;

#ifndef VARIANT
#define VARIANT 0
#endif

template <bool kFp8>
__global__ void __launch_bounds__(256, 1) bench_v0(int iters, float scale, long long* cyc, float* sink, uint32_t* pout) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const int t = warp >> 2;
  const uint32_t tS = tptr + ((uint32_t)((warp & 3) * 32) << 16) + t * 128;
  // fill S with small pseudo-random ints
  {
    uint32_t r[32];
#pragma unroll
    for (int c0 = 0; c0 < 128; c0 += 32) {
#pragma unroll
      for (int i = 0; i < 32; ++i) r[i] = (uint32_t)(int)(((threadIdx.x * 131 + (c0 + i) * 71) % 20011) - 10000);
      tmem_st32(tS + c0, r);
    }
    tc_wait_st();
  }
  float sc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) sc[k] = scale * (1.f + 0.01f * k);
  float m_run = -1e30f, d_run = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
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
    float alpha;
    uint32_t w[kFp8 ? 32 : 64];
    softmax_row<kFp8, false>(s, sc, 127, m_run, d_run, alpha, w);
    // P would overwrite S in the real kernel; here it goes to the O region so S stays intact
    if constexpr (kFp8) {
      tmem_st32(tS + 256, w);
    } else {
      uint32_t(&w0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[0]);
      uint32_t(&w1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[32]);
      tmem_st32(tS + 256, w0);
      tmem_st32(tS + 256 + 32, w1);
    }
    tc_wait_st();
    if (pout != nullptr && it == iters - 1 && blockIdx.x == 0) {
#pragma unroll
      for (int i = 0; i < (kFp8 ? 32 : 64); ++i) pout[threadIdx.x * (kFp8 ? 32 : 64) + i] = w[i];
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = d_run + m_run;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tptr);
}


// Streaming variant: stale-max chunks, packed arithmetic, kPN/kPD of the pairs on the polynomial.
template <bool kFp8, int kPN, int kPD, int kDeg>
__global__ void __launch_bounds__(256, 1) bench_stream(int iters, float scale, long long* cyc, float* sink,
                                                      uint32_t* pout) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const int t = warp >> 2;
  const uint32_t tS = tptr + ((uint32_t)((warp & 3) * 32) << 16) + t * 128;
  {
    uint32_t r[32];
#pragma unroll
    for (int c0 = 0; c0 < 128; c0 += 32) {
#pragma unroll
      for (int i = 0; i < 32; ++i) r[i] = (uint32_t)(int)(((threadIdx.x * 131 + (c0 + i) * 71) % 20011) - 10000);
      tmem_st32(tS + c0, r);
    }
    tc_wait_st();
  }
  SoftmaxConsts k;
#pragma unroll
  for (int i = 0; i < 8; ++i) k.sc[i] = scale * (1.f + 0.01f * i);
  float m_run = -1e30f, d_run = 0.f;
  constexpr float kTau = kFp8 ? kFp8Tau : kF16Tau;
  constexpr int kW = kFp8 ? 8 : 16;
  softmax_set_bias<kFp8>(k, m_run);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t w[4][kW];
    uint64_t dsum2;
    float alpha = 1.f;
    for (;;) {
      int mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = kMaskedS;
      dsum2 = 0;
      uint32_t a[32], b[32];
      tmem_ld32(tS, a);
      tc_wait_ld();
      tmem_ld32(tS + 32, b);
      softmax_chunk<kFp8, false, 0, kPN, kPD, kDeg>(a, k, 127, mx, dsum2, w[0]);
      tc_wait_ld();
      tmem_ld32(tS + 64, a);
      softmax_chunk<kFp8, false, 1, kPN, kPD, kDeg>(b, k, 127, mx, dsum2, w[1]);
      tc_wait_ld();
      tmem_ld32(tS + 96, b);
      softmax_chunk<kFp8, false, 2, kPN, kPD, kDeg>(a, k, 127, mx, dsum2, w[2]);
      tc_wait_ld();
      softmax_chunk<kFp8, false, 3, kPN, kPD, kDeg>(b, k, 127, mx, dsum2, w[3]);
      const float mrow = softmax_rowmax<false>(mx, k);
      if (!__any_sync(0xffffffffu, mrow > m_run + kTau)) break;
      const float m_new = fmaxf(m_run, mrow);
      alpha = ex2_approx(m_run - m_new);
      m_run = m_new;
      softmax_set_bias<kFp8>(k, m_run);
    }
    float d0, d1;
    unpack2(dsum2, d0, d1);
    d_run = fmaf(d_run, alpha, d0 + d1);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if constexpr (kFp8) tmem_st8(tS + 256 + c * kW, w[c]);
      else tmem_st16(tS + 256 + c * kW, w[c]);
    }
    tc_wait_st();
    if (pout != nullptr && it == iters - 1 && blockIdx.x == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < kW; ++i) pout[threadIdx.x * 4 * kW + c * kW + i] = w[c][i];
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = d_run + m_run;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tptr);
}

static int g_iters = 2000;
static float e4m3_to_float(uint8_t v) {
  const int e = (v >> 3) & 15, m = v & 7;
  const float f = e == 0 ? ldexpf((float)m, -9) : ldexpf(1.f + m / 8.f, e - 7);
  return (v & 0x80) ? -f : f;
}
static float half_to_float(uint16_t h) {
  const int e = (h >> 10) & 31, m = h & 1023;
  const float f = e == 0 ? ldexpf((float)m, -24) : ldexpf(1.f + m / 1024.f, e - 15);
  return (h & 0x8000) ? -f : f;
}

template <bool kFp8, int kPN, int kPD, int kDeg>
static void run_stream(const char* name, long long* cyc, float* sink, uint32_t* pout, const uint32_t* ref) {
  const int grid = 148, iters = g_iters;
  long long h[148];
  for (int rep = 0; rep < 2; ++rep) {
    bench_stream<kFp8, kPN, kPD, kDeg><<<grid, 256>>>(iters, 2e-5f, cyc, sink, pout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  // compare P with the baseline's
  const int words = 256 * (kFp8 ? 32 : 64);
  static uint32_t hp[256 * 64];
  cudaMemcpy(hp, pout, words * 4, cudaMemcpyDeviceToHost);
  double worst = 0, sum = 0;
  int ndiff = 0;
  for (int i = 0; i < words; ++i) {
    for (int b = 0; b < (kFp8 ? 4 : 2); ++b) {
      float x, y;
      if (kFp8) { x = e4m3_to_float(hp[i] >> (8 * b)); y = e4m3_to_float(ref[i] >> (8 * b)); }
      else { x = half_to_float(hp[i] >> (16 * b)); y = half_to_float(ref[i] >> (16 * b)); }
      const double d = fabs(x - y) / (fabs(y) + 1e-30);
      if (x != y) ++ndiff;
      worst = d > worst ? d : worst;
      sum += y;
    }
  }
  printf("%-28s %5.0f cycles per step pair -> %5.2f logits/clk/SM | vs row: %d of %d differ, worst rel %.2e (sum %.3g)\n",
         name, (double)mx / iters, 32768.0 * iters / mx, ndiff, words * (kFp8 ? 4 : 2), worst, sum);
}
// Warp-count scaling: kThreads/32 warps per SM, each thread handles kChunks 32-column chunks of one
// row (4 = whole row, 2 = half row).  Timing only (several warps may share an S tile).
template <bool kFp8, int kThreads, int kChunks, int kPN, int kPD, int kDeg>
__global__ void __launch_bounds__(kThreads, 1) bench_warps(int iters, float scale, long long* cyc, float* sink) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const int grp = warp >> 2;
  const uint32_t tS = tptr + ((uint32_t)((warp & 3) * 32) << 16) + (grp & 1) * 128 + ((grp >> 1) & 1) * (kChunks == 2 ? 64 : 0);
  if (grp < 2) {
    uint32_t r[32];
#pragma unroll
    for (int c0 = 0; c0 < 128; c0 += 32) {
#pragma unroll
      for (int i = 0; i < 32; ++i) r[i] = (uint32_t)(int)(((threadIdx.x * 131 + (c0 + i) * 71) % 20011) - 10000);
      tmem_st32(tS + c0, r);
    }
    tc_wait_st();
  }
  SoftmaxConsts k;
#pragma unroll
  for (int i = 0; i < 8; ++i) k.sc[i] = scale * (1.f + 0.01f * i);
  float m_run = -1e30f, d_run = 0.f;
  constexpr float kTau = kFp8 ? kFp8Tau : kF16Tau;
  constexpr int kW = kFp8 ? 8 : 16;
  softmax_set_bias<kFp8>(k, m_run);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t w[kChunks][kW];
    uint64_t dsum2;
    float alpha = 1.f;
    for (;;) {
      int mx[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = kMaskedS;
      dsum2 = 0;
      uint32_t a[32], b[32];
      tmem_ld32(tS, a);
      tc_wait_ld();
      tmem_ld32(tS + 32, b);
      softmax_chunk<kFp8, false, 0, kPN, kPD, kDeg>(a, k, 127, mx, dsum2, w[0]);
      tc_wait_ld();
      if constexpr (kChunks == 4) tmem_ld32(tS + 64, a);
      softmax_chunk<kFp8, false, 1, kPN, kPD, kDeg>(b, k, 127, mx, dsum2, w[1]);
      if constexpr (kChunks == 4) {
        tc_wait_ld();
        tmem_ld32(tS + 96, b);
        softmax_chunk<kFp8, false, 2, kPN, kPD, kDeg>(a, k, 127, mx, dsum2, w[2]);
        tc_wait_ld();
        softmax_chunk<kFp8, false, 3, kPN, kPD, kDeg>(b, k, 127, mx, dsum2, w[3]);
      }
      const float mrow = softmax_rowmax<false>(mx, k);
      if (!__any_sync(0xffffffffu, mrow > m_run + kTau)) break;
      const float m_new = fmaxf(m_run, mrow);
      alpha = ex2_approx(m_run - m_new);
      m_run = m_new;
      softmax_set_bias<kFp8>(k, m_run);
    }
    float d0, d1;
    unpack2(dsum2, d0, d1);
    d_run = fmaf(d_run, alpha, d0 + d1);
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
      if constexpr (kFp8) tmem_st8(tptr + ((uint32_t)((warp & 3) * 32) << 16) + 256 + grp * 32 + c * kW, w[c]);
      else tmem_st16(tptr + ((uint32_t)((warp & 3) * 32) << 16) + 256 + grp * 64 + c * kW, w[c]);
    }
    tc_wait_st();
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = d_run + m_run;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tptr);
}

template <bool kFp8, int kThreads, int kChunks, int kPN, int kPD, int kDeg>
static void run_warps(const char* name, long long* cyc, float* sink) {
  const int grid = 148, iters = g_iters;
  long long h[148];
  for (int rep = 0; rep < 2; ++rep) {
    bench_warps<kFp8, kThreads, kChunks, kPN, kPD, kDeg><<<grid, kThreads>>>(iters, 2e-5f, cyc, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  const double logits = (double)kThreads * kChunks * 32 * iters;
  printf("%-44s %5.0f cycles per 32768 logits -> %5.2f logits/clk/SM\n", name, mx * 32768.0 / logits, logits / mx);
}

int main(int argc, char** argv) {
  if (argc > 1) g_iters = atoi(argv[1]);
  const int grid = 148, iters = g_iters;
  long long* cyc;
  float* sink;
  uint32_t *pout, *pref;
  cudaMalloc(&cyc, grid * 8);
  cudaMalloc(&sink, grid * 1024 * 4);
  cudaMalloc(&pout, 256 * 64 * 4);
  cudaMalloc(&pref, 256 * 64 * 4);
  static uint32_t ref8[256 * 32], ref16[256 * 64];
  long long h[148];
  for (int fp8 = 1; fp8 >= 0; --fp8) {
    for (int rep = 0; rep < 2; ++rep) {
      if (fp8) bench_v0<true><<<grid, 256>>>(iters, 2e-5f, cyc, sink, pref);
      else bench_v0<false><<<grid, 256>>>(iters, 2e-5f, cyc, sink, pref);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    cudaMemcpy(fp8 ? ref8 : ref16, pref, 256 * (fp8 ? 32 : 64) * 4, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("softmax_row<%s> (production)  %5.0f cycles per step pair -> %5.2f logits/clk/SM\n",
           fp8 ? "fp8" : "f16", (double)mx / iters, 32768.0 * iters / mx);
  }
  run_stream<true, 0, 1, 2>("stream fp8 poly 0", cyc, sink, pout, ref8);
  run_stream<true, 1, 4, 2>("stream fp8 poly 1/4 deg2", cyc, sink, pout, ref8);
  run_stream<true, 1, 3, 2>("stream fp8 poly 1/3 deg2", cyc, sink, pout, ref8);
  run_stream<true, 3, 8, 2>("stream fp8 poly 3/8 deg2", cyc, sink, pout, ref8);
  run_stream<true, 1, 2, 2>("stream fp8 poly 1/2 deg2", cyc, sink, pout, ref8);
  run_stream<true, 1, 3, 3>("stream fp8 poly 1/3 deg3", cyc, sink, pout, ref8);
  run_stream<false, 0, 1, 3>("stream f16 poly 0", cyc, sink, pout, ref16);
  run_stream<false, 1, 4, 3>("stream f16 poly 1/4 deg3", cyc, sink, pout, ref16);
  run_stream<false, 1, 3, 3>("stream f16 poly 1/3 deg3", cyc, sink, pout, ref16);
  run_stream<false, 3, 8, 3>("stream f16 poly 3/8 deg3", cyc, sink, pout, ref16);
  run_stream<false, 1, 2, 3>("stream f16 poly 1/2 deg3", cyc, sink, pout, ref16);
  run_warps<true, 128, 4, 0, 1, 2>("fp8 1 warp/sched full row poly 0", cyc, sink);
  run_warps<true, 256, 4, 0, 1, 2>("fp8 2 warps/sched full row poly 0", cyc, sink);
  run_warps<true, 512, 4, 0, 1, 2>("fp8 4 warps/sched full row poly 0", cyc, sink);
  run_warps<true, 512, 2, 0, 1, 2>("fp8 4 warps/sched half row poly 0", cyc, sink);
  run_warps<true, 128, 4, 1, 4, 2>("fp8 1 warp/sched full row poly 1/4", cyc, sink);
  run_warps<true, 512, 4, 1, 4, 2>("fp8 4 warps/sched full row poly 1/4", cyc, sink);
  run_warps<true, 512, 2, 1, 4, 2>("fp8 4 warps/sched half row poly 1/4", cyc, sink);
  run_warps<true, 512, 2, 3, 8, 2>("fp8 4 warps/sched half row poly 3/8", cyc, sink);
  run_warps<true, 512, 2, 1, 2, 2>("fp8 4 warps/sched half row poly 1/2", cyc, sink);
  run_warps<false, 128, 4, 0, 1, 3>("f16 1 warp/sched full row poly 0", cyc, sink);
  run_warps<false, 512, 2, 0, 1, 3>("f16 4 warps/sched half row poly 0", cyc, sink);
  run_warps<false, 512, 2, 1, 4, 3>("f16 4 warps/sched half row poly 1/4", cyc, sink);
  run_warps<false, 512, 2, 3, 8, 3>("f16 4 warps/sched half row poly 3/8", cyc, sink);
  run_warps<false, 512, 2, 1, 2, 3>("f16 4 warps/sched half row poly 1/2", cyc, sink);
  return 0;
}
