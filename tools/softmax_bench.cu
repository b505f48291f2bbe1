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
#include "softmax_sm100.cuh"

using namespace sage;

#ifndef VARIANT
#define VARIANT 0
#endif

template <bool kFp8>
__global__ void __launch_bounds__(256, 1) bench_v0(int iters, float scale, long long* cyc, float* sink) {
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
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = d_run + m_run;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tptr);
}

int main() {
  const int grid = 148, iters = 2000;
  long long* cyc;
  float* sink;
  cudaMalloc(&cyc, grid * 8);
  cudaMalloc(&sink, grid * 1024 * 4);
  long long h[148];
  for (int fp8 = 1; fp8 >= 0; --fp8) {
    for (int rep = 0; rep < 2; ++rep) {
      if (fp8) bench_v0<true><<<grid, 256>>>(iters, 2e-5f, cyc, sink);
      else bench_v0<false><<<grid, 256>>>(iters, 2e-5f, cyc, sink);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("softmax_row<%s> 8 warps/SM, 128 logits/thread/step: %.0f cycles per step pair (32768 logits) -> %.2f logits/clk/SM\n",
           fp8 ? "fp8" : "f16", (double)mx / iters, 32768.0 * iters / mx);
  }
  return 0;
}
