// How fast can 1 / 2 / 4 warps per SM sub-partition run the exp phase alone (registers only)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Isageattention_b200/csrc -o tools/_bin/exp_bench tools/exp_bench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "softmax_sm100.cuh"
using namespace sage;

template <int kThreads, int kMode>
__global__ void __launch_bounds__(kThreads, 1) k_exp(int iters, float scale, long long* cyc, float* sink) {
  uint32_t s[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) s[i] = (uint32_t)(int)(((threadIdx.x * 131 + i * 71) % 20011) - 10000);
  float sc[8], bias[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sc[k] = scale * (1.f + 0.01f * k); bias[k] = fmaf(-kMagicF, sc[k], -0.1f * k); }
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    acc += softmax_exp<false>(s, sc, bias, 127);
    if (kMode == 1) {
      uint32_t w[32];
      softmax_pack<true>(s, w);
#pragma unroll
      for (int i = 0; i < 32; ++i) { s[4 * i] = w[i] & 0x3fff; s[4 * i + 1] = (w[i] >> 8) & 0x3fff; s[4 * i + 2] = (w[i] >> 16) & 0x3fff; s[4*i+3] = w[i] >> 18; }
    } else {
#pragma unroll
      for (int i = 0; i < 128; ++i) s[i] = s[i] & 0x3fff;  // p bits -> small positive ints again (1 LOP per logit)
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int kThreads, int kMode>
void run(const char* name, long long* cyc, float* sink) {
  const int iters = 1000;
  long long h[148];
  for (int r = 0; r < 2; ++r) {
    k_exp<kThreads, kMode><<<148, kThreads>>>(iters, 2e-5f, cyc, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  const double per_warp_row = (double)mx / iters;                       // cycles per 128-logit row per warp
  const double per_mufu_sched = per_warp_row / 128.0 / (kThreads / 128);  // cycles per MUFU warp-instr per scheduler
  printf("%-34s %7.0f cycles per row-step per warp, %5.2f cycles per MUFU per scheduler\n", name, per_warp_row, per_mufu_sched);
}

int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 4);
  run<128, 0>("exp only, 1 warp/sched", cyc, sink);
  run<256, 0>("exp only, 2 warps/sched", cyc, sink);
  run<512, 0>("exp only, 4 warps/sched", cyc, sink);
  run<128, 1>("exp+pack e4m3, 1 warp/sched", cyc, sink);
  run<256, 1>("exp+pack e4m3, 2 warps/sched", cyc, sink);
  return 0;
}
