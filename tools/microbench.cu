// tools/microbench.cu — sm_100a issue/pipe micro-benchmarks that the filter-loop design rests on.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Prints, per variant, FP32 FMA lanes per clock per SM (128 = the FP32 pipe's peak) from
// clock64() on one resident wave, and TFLOP/s from CUDA events.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b),
                     rc = *reinterpret_cast<unsigned long long *>(&c), rd;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2 *>(&rd);
}

// MODE 0: 16 scalar FFMA chains.  1: 8 FFMA2 chains.  2: FFMA2 + 2 FSETP per 7.  3: + LDS.128 per 7.
// 4: scalar FFMA + 1 FSETP per 7 (the round-0 loop shape).  5: FFMA2 + 1 FMNMX3-ish (2 FMNMX) per 7.
template <int MODE>
__global__ void __launch_bounds__(256) bench(float *out, int iters, float bb, float cc, long long *cyc) {
  __shared__ float4 sm[64];
  if (threadIdx.x < 64) sm[threadIdx.x] = make_float4(bb, cc, bb, cc);
  __syncthreads();
  float2 a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = make_float2((float)(threadIdx.x + k), (float)k);
  float2 b = make_float2(bb, bb), c = make_float2(cc, cc);
  float thr = 1e30f * bb;
  int cnt = 0;
  float mx = 0.f;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (MODE == 3) {
        float4 s = sm[(i + u) & 63];
        b = make_float2(s.x, s.y); c = make_float2(s.z, s.w);
      }
      if (MODE == 0 || MODE == 4) {
#pragma unroll
        for (int k = 0; k < 7; ++k) { a[k].x = fmaf(a[k].x, b.x, c.x); a[k].y = fmaf(a[k].y, b.y, c.y); }
        if (MODE == 4) { if (a[0].x >= thr) cnt++; if (a[0].y >= thr) cnt++; }
      } else {
#pragma unroll
        for (int k = 0; k < 7; ++k) a[k] = ffma2(a[k], b, c);
        if (MODE == 2 || MODE == 3) { if (a[0].x >= thr) cnt++; if (a[0].y >= thr) cnt++; }
        if (MODE == 5) mx = fmaxf(mx, fmaxf(a[0].x, a[0].y));
      }
    }
  }
  long long t1 = clock64();
  float s = (float)cnt + mx;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k].x + a[k].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int sms, int blocks_per_sm) {
  int blocks = sms * blocks_per_sm, iters = 20000;
  float *out; long long *cyc;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  CK(cudaMalloc(&cyc, blocks * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    bench<MODE><<<blocks, 256>>>(out, iters, 0.999f, 0.001f, cyc);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  long long *h = (long long *)malloc(blocks * 8);
  CK(cudaMemcpy(h, cyc, blocks * 8, cudaMemcpyDeviceToHost));
  double avg = 0; for (int k = 0; k < blocks; ++k) avg += (double)h[k]; avg /= blocks;
  double fma_per_thread = 14.0 * 4 * iters;  // 7 pairs x 2 lanes x 4 unrolled groups
  double lanes_per_clk_sm = fma_per_thread * 256 * blocks_per_sm / avg;
  double tflops = 2.0 * fma_per_thread * 256 * blocks / (best * 1e-3) / 1e12;
  printf("%-34s %2d blk/SM  %7.2f FMA-lanes/clk/SM  %7.2f TFLOP/s  (%.3f ms, %.0f cyc, %.0f MHz)\n", name,
         blocks_per_sm, lanes_per_clk_sm, tflops, best, avg, avg / (best * 1e3));
  cudaFree(out); cudaFree(cyc); free(h);
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("%s, %d SMs\n", p.name, sms);
  for (int bps : {2, 4, 8}) {
    run<0>("FFMA x14", sms, bps);
    run<1>("FFMA2 x7", sms, bps);
    run<4>("FFMA x14 + 2 FSETP", sms, bps);
    run<2>("FFMA2 x7 + 2 FSETP", sms, bps);
    run<3>("FFMA2 x7 + 2 FSETP + LDS.128", sms, bps);
    run<5>("FFMA2 x7 + 2 FMNMX", sms, bps);
  }
  return 0;
}
