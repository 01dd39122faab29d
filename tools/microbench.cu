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


// ---- filter-loop replicas: the path kernel's filter without the rare path.  NP slot pairs,
// G rows per group; rows from shared memory (broadcast LDS.128 -> .F32 register operands) or
// from the constant bank through uniform registers (LDCU -> .F32 uniform operands).
__device__ __forceinline__ float2 bc(float v) { return make_float2(v, v); }
struct Rows { float4 row[520]; };

template <int NP, int G, bool CONSTANT>
__global__ void __launch_bounds__(256) feed(float *out, int iters, const float4 *g_rows, float thr,
                                            const __grid_constant__ Rows CR) {
  __shared__ float4 s_rows[520];
  for (int k = threadIdx.x; k < 520; k += 256) s_rows[k] = g_rows[k];
  __syncthreads();
  float2 D0[NP], D1[NP], D2[NP], N0[NP], O0[NP], O1[NP], O2[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const float *g = reinterpret_cast<const float *>(g_rows) + threadIdx.x + 16 * p;
    D0[p] = make_float2(g[0], g[1]); D1[p] = make_float2(g[2], g[3]); D2[p] = make_float2(g[4], g[5]);
    N0[p] = make_float2(g[6], g[7]); O0[p] = make_float2(g[8], g[9]); O1[p] = make_float2(g[10], g[11]);
    O2[p] = make_float2(g[12], g[13]);
  }
  float sink = 0.f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll 2
    for (int s0 = 0; s0 < 512; s0 += G) {
      float mx[NP], my[NP];
#pragma unroll
      for (int k = 0; k < G; ++k) {
        float4 S = CONSTANT ? CR.row[s0 + k] : s_rows[s0 + k];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          float2 b = ffma2(bc(S.x), D0[p], ffma2(bc(S.y), D1[p], ffma2(bc(S.z), D2[p], N0[p])));
          float2 e1 = ffma2(b, b, bc(S.w));
          float2 e = ffma2(bc(S.x), O0[p], ffma2(bc(S.y), O1[p], ffma2(bc(S.z), O2[p], e1)));
          mx[p] = k ? fmaxf(mx[p], e.x) : e.x;
          my[p] = k ? fmaxf(my[p], e.y) : e.y;
        }
      }
      bool any = false;
#pragma unroll
      for (int p = 0; p < NP; ++p) any = any || mx[p] >= thr || my[p] >= thr;
      if (__any_sync(0xffffffffu, any)) sink += 1.f;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = sink;
}

template <int NP, int G, bool CONSTANT>
void run_feed(const char *name, int sms, int blocks_per_sm) {
  int blocks = sms * blocks_per_sm, iters = 200;
  float *out; float4 *rows;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  CK(cudaMalloc(&rows, 520 * 16));
  static Rows h;
  for (int k = 0; k < 520; ++k) h.row[k] = make_float4(k * 1e-3f, 1.f - k * 1e-3f, 0.5f + k * 1e-4f, -1e3f - k);
  CK(cudaMemcpy(rows, h.row, sizeof h, cudaMemcpyHostToDevice));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    feed<NP, G, CONSTANT><<<blocks, 256>>>(out, iters, rows, 1e30f, h);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  double fma = 14.0 * NP * 512.0 * iters * 256.0 * blocks;
  printf("%-40s %2d blk/SM  %7.2f TFLOP/s  (%.3f ms)\n", name, blocks_per_sm, 2.0 * fma / (best * 1e-3) / 1e12, best);
  cudaFree(out); cudaFree(rows);
}


// ---- register-file pressure: every FMA reads three DISTINCT register operands per lane.
// V=0: scalar FFMA d=a*b+c (3 reads).  V=1: FFMA2 with three pair operands (6 reads / 2 FMA).
// V=2: FFMA2 with a scalar-broadcast multiplier (.F32) (5 reads).  V=3: FFMA2 with the
// multiplier in a uniform register (4 reads).
template <int V>
__global__ void __launch_bounds__(256) rfbench(float *out, int iters, const float *g, float u0) {
  float2 a[8], b[8], c[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = make_float2(g[threadIdx.x + k], g[threadIdx.x + k + 8]);
    b[k] = make_float2(g[threadIdx.x + k + 16], g[threadIdx.x + k + 24]);
    c[k] = make_float2(g[threadIdx.x + k + 32], g[threadIdx.x + k + 40]);
  }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int k1 = (k + 1 + u) & 7, k2 = (k + 3 + u) & 7;
        if (V == 0) { a[k].x = fmaf(b[k1].x, c[k2].x, a[k].x); a[k].y = fmaf(b[k1].y, c[k2].y, a[k].y); }
        if (V == 1) a[k] = ffma2(b[k1], c[k2], a[k]);
        if (V == 2) a[k] = ffma2(bc(b[k1].x), c[k2], a[k]);
        if (V == 3) a[k] = ffma2(bc(u0), c[k2], a[k]);
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k].x + a[k].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int V>
void run_rf(const char *name, int sms, int blocks_per_sm) {
  int blocks = sms * blocks_per_sm, iters = 5000;
  float *out, *g;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  CK(cudaMalloc(&g, 1024 * 4));
  CK(cudaMemset(g, 0, 1024 * 4));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    rfbench<V><<<blocks, 256>>>(out, iters, g, 0.5f);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  double fma = 16.0 * 4 * iters * 256.0 * blocks;
  printf("%-44s %2d blk/SM  %7.2f TFLOP/s  (%.3f ms)\n", name, blocks_per_sm, 2.0 * fma / (best * 1e-3) / 1e12, best);
  cudaFree(out); cudaFree(g);
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("%s, %d SMs\n", p.name, sms);
  for (int bps : {8}) {
    run<0>("FFMA x14", sms, bps);
    run<1>("FFMA2 x7", sms, bps);
    run<4>("FFMA x14 + 2 FSETP", sms, bps);
    run<2>("FFMA2 x7 + 2 FSETP", sms, bps);
    run<3>("FFMA2 x7 + 2 FSETP + LDS.128", sms, bps);
    run<5>("FFMA2 x7 + 2 FMNMX", sms, bps);
  }
  for (int bps : {2, 4}) {
    run_rf<0>("FFMA  3 distinct reg operands", sms, bps);
    run_rf<1>("FFMA2 3 distinct pair operands", sms, bps);
    run_rf<2>("FFMA2 scalar-broadcast multiplier (.F32 R)", sms, bps);
    run_rf<3>("FFMA2 uniform-register multiplier (.F32 UR)", sms, bps);
  }
  for (int bps : {2, 4}) {
    run_feed<1, 4, false>("filter R=2 G=4 rows via LDS.128", sms, bps);
    run_feed<1, 4, true>("filter R=2 G=4 rows via LDCU", sms, bps);
    run_feed<1, 8, true>("filter R=2 G=8 rows via LDCU", sms, bps);
    run_feed<2, 4, false>("filter R=4 G=4 rows via LDS.128", sms, bps);
    run_feed<2, 4, true>("filter R=4 G=4 rows via LDCU", sms, bps);
    run_feed<2, 2, true>("filter R=4 G=2 rows via LDCU", sms, bps);
  }
  return 0;
}
