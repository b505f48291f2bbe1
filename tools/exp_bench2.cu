// Interference test: warps 0-3 (one per SM sub-partition) run the exp phase in a loop and are
// timed; warps 4-7 run a companion loop of one other softmax-step ingredient until the timed
// warps finish.  Which companion slows the exp warps down?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Isageattention_b200/csrc -o tools/_bin/exp_bench2 tools/exp_bench2.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "softmax_sm100.cuh"
using namespace sage;

enum { kNone = 0, kExp, kMax, kPackF8, kPackF16, kSts, kLdtm, kSttm, kMbar, kFence };

template <int kComp>
__global__ void __launch_bounds__(256, 1) k_mix(int iters, float scale, long long* cyc, float* sink) {
  __shared__ uint32_t tptr;
  __shared__ volatile int done;
  __shared__ uint64_t bar;
  __shared__ __align__(16) uint32_t buf[128 * 32];
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { done = 0; mbar_init(&bar, 1); }
  if (warp == 0) tmem_alloc<512>(&tptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t s[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) s[i] = (uint32_t)(int)(((threadIdx.x * 131 + i * 71) % 20011) - 10000);
  float sc[8], bias[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sc[k] = scale * (1.f + 0.01f * k); bias[k] = fmaf(-kMagicF, sc[k], -0.1f * k); }
  float acc = 0.f;
  const uint32_t taddr = tptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
  __syncthreads();
  if (warp < 4 || kComp == kExp) {
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      acc += softmax_exp<false>(s, sc, bias, 127);
#pragma unroll
      for (int i = 0; i < 128; ++i) s[i] = s[i] & 0x3fff;
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    __syncwarp();
    if (warp < 4 && (threadIdx.x & 31) == 0) atomicAdd((int*)&done, 1);
  } else {
    uint32_t w[64];
    while (done < 4) {
      if constexpr (kComp == kMax) {
        float a; softmax_max<true, false>(s, sc, 127, acc, a, bias);
#pragma unroll
        for (int i = 0; i < 128; ++i) s[i] += 3;
      } else if constexpr (kComp == kPackF8) {
        uint32_t(&w8)[32] = *reinterpret_cast<uint32_t(*)[32]>(&w[0]);
        softmax_pack<true>(s, w8);
#pragma unroll
        for (int i = 0; i < 32; ++i) s[4 * i] ^= w8[i];
      } else if constexpr (kComp == kPackF16) {
        softmax_pack<false>(s, w);
#pragma unroll
        for (int i = 0; i < 64; ++i) s[2 * i] ^= w[i];
      } else if constexpr (kComp == kSts) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
          st_shared_v4(smem_u32(buf) + (threadIdx.x & 127) * 128 + ((c ^ (threadIdx.x & 7)) << 4), s[4 * c], s[4 * c + 1], s[4 * c + 2], s[4 * c + 3]);
      } else if constexpr (kComp == kFence) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
          st_shared_v4(smem_u32(buf) + (threadIdx.x & 127) * 128 + ((c ^ (threadIdx.x & 7)) << 4), s[4 * c], s[4 * c + 1], s[4 * c + 2], s[4 * c + 3]);
        fence_proxy_async_smem();
      } else if constexpr (kComp == kLdtm) {
        uint32_t(&a)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&b)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        tmem_ld32(taddr, a); tmem_ld32(taddr + 32, b); tc_wait_ld();
      } else if constexpr (kComp == kSttm) {
        uint32_t(&a)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        tmem_st32(taddr, a); tc_wait_st();
      } else if constexpr (kComp == kMbar) {
        acc += mbar_try_wait(&bar, 0, 100) ? 1.f : 0.f;
      }
    }
#pragma unroll
    for (int i = 0; i < 128; ++i) acc += __uint_as_float(s[i]);
  }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tptr);
}

template <int kComp>
void run(const char* name, long long* cyc, float* sink) {
  const int iters = 500;
  long long h[148];
  for (int r = 0; r < 2; ++r) {
    k_mix<kComp><<<148, 256>>>(iters, 2e-5f, cyc, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("companion %-28s exp warp: %6.0f cycles per 128-logit row\n", name, (double)mx / iters);
}

int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 4);
  run<kNone>("idle", cyc, sink);
  run<kExp>("exp (2 exp warps/sched)", cyc, sink);
  run<kMax>("row max (VIMNMX3)", cyc, sink);
  run<kPackF8>("pack e4m3 (F2FP+PRMT)", cyc, sink);
  run<kPackF16>("pack f16 (F2FP)", cyc, sink);
  run<kSts>("st.shared.v4 x8", cyc, sink);
  run<kFence>("st.shared.v4 x8 + proxy fence", cyc, sink);
  run<kLdtm>("tcgen05.ld x2 + wait", cyc, sink);
  run<kSttm>("tcgen05.st + wait", cyc, sink);
  run<kMbar>("mbarrier try_wait spin", cyc, sink);
  return 0;
}
