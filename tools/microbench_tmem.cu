// tools/microbench_tmem.cu — how fast does tensor memory come back to registers on this GPU?
// Every warp of a CTA (one CTA per SM, 512 TMEM columns) reads its lane quarter of a 64-column
// window over and over with tcgen05.ld.32x32b.x32 (x2), like the read-back of filter_rays_tc, and the
// kernel reports bytes per clock and SM for 4 .. 24 warps.   nvcc -arch=sm_100a -o mb_tmem tools/microbench_tmem.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define LD32(addr, v)                                                                                        \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, " \
               "%12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, "  \
               "%30, %31}, [%32];"                                                                          \
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),        \
                 "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),    \
                 "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), \
                 "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), \
                 "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                                         \
               : "r"(addr)                                                                                  \
               : "memory")

__global__ void __launch_bounds__(768, 1) tmem_read(int iters, unsigned *sink, long long *cycles) {
  __shared__ uint32_t s_tmem;
  const uint32_t warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // the warp's lane quarter, a 64-column window per warpgroup
  const uint32_t addr = s_tmem + (((warp & 3u) * 32u) << 16) + ((warp >> 2) % 8u) * 64u;
  unsigned acc = 0;
  uint32_t v[32];
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    LD32(addr, v);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    acc ^= v[0] ^ v[13] ^ v[31];
    LD32(addr + 32u, v);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    acc ^= v[1] ^ v[17] ^ v[30];
  }
  const long long t1 = clock64();
  if (acc == 0x12345678u) sink[0] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(s_tmem) : "memory");
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  unsigned *sink; long long *cyc;
  cudaMalloc(&sink, 4); cudaMalloc(&cyc, 8 * p.multiProcessorCount);
  const int iters = 20000;
  for (int warps : {4, 8, 12, 16, 20, 24}) {
    tmem_read<<<p.multiProcessorCount, warps * 32, 0>>>(100, sink, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    tmem_read<<<p.multiProcessorCount, warps * 32, 0>>>(iters, sink, cyc);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double bytes = (double)iters * 2 * 32 * 32 * 4 * warps;  // per SM
    printf("warps %2d: %8.3f ms, %lld cycles (SM 0), %.1f B/clk/SM, %.1f GB/s/SM  %s\n", warps, ms, c, bytes / (double)c,
           bytes / (ms * 1e6), e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
