// Pipe-throughput microbenchmarks for the softmax inner loop on sm_100a (development tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/microbench tools/microbench.cu
// Prints results per SM per clock for each op, with 8 and 16 resident warps per SM.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define ITERS 2048
#define ILP 8

template <int OP>
__device__ __forceinline__ void body(uint32_t (&r)[ILP], uint32_t c0, uint32_t c1) {
#pragma unroll
  for (int i = 0; i < ILP; ++i) {
    if constexpr (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(r[i]));
    if constexpr (OP == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r[i]));
    if constexpr (OP == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r[i]));
    if constexpr (OP == 3) {
      uint16_t h;
      asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(h) : "f"(__uint_as_float(r[i])), "f"(__uint_as_float(c0)));
      r[i] ^= h;
    }
    if constexpr (OP == 4) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(__uint_as_float(r[i])), "f"(__uint_as_float(c0)));
    if constexpr (OP == 5) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c0), "r"(c1));
    if constexpr (OP == 7) asm volatile("max.s32 %0, %0, %1;" : "+r"(r[i]) : "r"(c0 + i));
    if constexpr (OP == 8) asm volatile("max.f32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c0), "r"(c1));
    if constexpr (OP == 9) asm volatile("cvt.rn.f32.s32 %0, %0;" : "+r"(r[i]));
    if constexpr (OP == 10) {
      uint16_t h;
      asm volatile("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(h) : "r"(r[i]));
      r[i] ^= h;
    }
    if constexpr (OP == 11) asm volatile("add.s32 %0, %0, %1;" : "+r"(r[i]) : "r"(c0));
    if constexpr (OP == 12) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c0), "r"(c1));
    if constexpr (OP == 13) asm volatile("add.rn.f32 %0, %0, %1;" : "+r"(r[i]) : "r"(c0));
    if constexpr (OP == 14) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(c0), "r"(c1));
    if constexpr (OP == 15) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c0), "r"(c1));
  }
}

// packed f32x2 ops use 64-bit registers
template <int OP>
__device__ __forceinline__ void body2(uint64_t (&r)[ILP], uint64_t c0, uint64_t c1) {
#pragma unroll
  for (int i = 0; i < ILP; ++i) {
    if constexpr (OP == 0) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(r[i]) : "l"(c0), "l"(c1));
    if constexpr (OP == 1) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(r[i]) : "l"(c0));
    if constexpr (OP == 2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(r[i]) : "l"(c0));
  }
}

template <int OP>
__global__ void k32(uint32_t* out, long long* cyc, uint32_t c0, uint32_t c1) {
  uint32_t r[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) r[i] = threadIdx.x * 7 + i + c0;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) body<OP>(r, c0, c1);
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc ^= r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP>
__global__ void k64(uint64_t* out, long long* cyc, uint64_t c0, uint64_t c1) {
  uint64_t r[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) r[i] = threadIdx.x * 7 + i + c0;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) body2<OP>(r, c0, c1);
  long long t1 = clock64();
  uint64_t acc = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc ^= r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// TMEM load bandwidth: each warp of the first 4 (or 8) warps reads its 32-lane quarter, x32 columns per op
template <int NCOL>
__global__ void ktmem(uint32_t* out, long long* cyc) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < 256; ++it) {
#pragma unroll
    for (int c = 0; c < 512; c += NCOL) {
      if constexpr (NCOL == 32) {
        uint32_t r[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
              "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(base + c)
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= r[i];
      }
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tptr) : "memory");
}

static long long maxcyc(long long* d, int n) {
  long long h[4096];
  cudaMemcpy(h, d, n * sizeof(long long), cudaMemcpyDeviceToHost);
  long long m = 0;
  for (int i = 0; i < n; ++i) m = h[i] > m ? h[i] : m;
  return m;
}

#define RUN32(OP, NAME, ELEMS)                                                          \
  for (int thr : {256, 512}) {                                                          \
    k32<OP><<<grid, thr>>>(out, cyc, 0x3f000000u, 0x3e000000u);                         \
    k32<OP><<<grid, thr>>>(out, cyc, 0x3f000000u, 0x3e000000u);                         \
    cudaDeviceSynchronize();                                                            \
    long long c = maxcyc(cyc, grid);                                                    \
    printf("%-34s warps/SM=%2d  %7.2f elem/clk/SM  (%6.2f instr/clk/SM)\n", NAME, thr / 32, \
           (double)ITERS * ILP * thr * ELEMS / c, (double)ITERS * ILP * thr / 32 / c);   \
  }
#define RUN64(OP, NAME)                                                                 \
  for (int thr : {256, 512}) {                                                          \
    k64<OP><<<grid, thr>>>((uint64_t*)out, cyc, 0x3f0000003f000000ull, 0x3e0000003e000000ull); \
    k64<OP><<<grid, thr>>>((uint64_t*)out, cyc, 0x3f0000003f000000ull, 0x3e0000003e000000ull); \
    cudaDeviceSynchronize();                                                            \
    long long c = maxcyc(cyc, grid);                                                    \
    printf("%-34s warps/SM=%2d  %7.2f elem/clk/SM  (%6.2f instr/clk/SM)\n", NAME, thr / 32, \
           (double)ITERS * ILP * thr * 2 / c, (double)ITERS * ILP * thr / 32 / c);       \
  }

int main() {
  int grid = 148;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 512 * 8);
  cudaMalloc(&cyc, 4096 * 8);
  RUN32(0, "ex2.approx.ftz.f32", 1)
  RUN32(1, "ex2.approx.ftz.f16x2", 2)
  RUN32(2, "ex2.approx.ftz.bf16x2", 2)
  RUN32(3, "cvt.rn.satfinite.e4m3x2.f32", 2)
  RUN32(4, "cvt.rn.f16x2.f32", 2)
  RUN32(10, "cvt.rn.satfinite.e4m3x2.f16x2", 2)
  RUN32(5, "fma.rn.f32", 1)
  RUN64(0, "fma.rn.f32x2")
  RUN64(1, "add.rn.f32x2")
  RUN64(2, "mul.rn.f32x2")
  RUN32(13, "add.rn.f32", 1)
  RUN32(12, "fma.rn.f16x2", 2)
  RUN32(7, "max.s32", 1)
  RUN32(8, "max.f32 (3-input)", 2)
  RUN32(9, "cvt.rn.f32.s32", 1)
  RUN32(11, "add.s32", 1)
  RUN32(14, "lop3.b32", 1)
  RUN32(15, "prmt.b32", 1)
  for (int thr : {128, 256}) {
    ktmem<32><<<grid, thr>>>(out, cyc);
    ktmem<32><<<grid, thr>>>(out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = maxcyc(cyc, grid);
    double bytes = 256.0 * 512 * 4 * thr;  // per CTA
    printf("tcgen05.ld 32x32b.x32 + wait        warps/SM=%2d  %7.1f B/clk/SM  (%s)\n", thr / 32, bytes / c,
           cudaGetErrorString(e));
  }
  return 0;
}
