// tools/microbench_issue.cu — what does an instruction cost when it is issued next to a stream of
// FFMA2 (sm_100a)?  FFMA2 holds the FMA pipe for two cycles; the question the filter loop's shape
// rests on is which other instructions can be issued in the second cycle ("the shadow") and which
// cost a slot of their own.  Per variant: 28 FFMA2 (4 independent chains of 7, multiplier in a
// register pair) plus N instructions of one kind, interleaved; prints cycles per iteration per warp
// on one scheduler (CUDA-event time of one resident wave x the SM clock, W warps per scheduler) --
// 56.0 = the FFMA2s alone.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench_issue tools/microbench_issue.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

typedef unsigned long long u64;
__device__ __forceinline__ void ffma2(u64 &a, u64 b, u64 c) {
  asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a) : "l"(b), "l"(c));
}

struct Rows { float4 row[256]; };  // 8 x 28 rows used

enum Kind { NONE, FMNMX, FSETP, ISETP, IADD, LOP, IMAD, VOTE, LDCU_FEED, FMUL, MOVK, SHFL, LDS, FADD2, UBRA };

// K = kind, N = extra instructions per 28 FFMA2 (a multiple of 4)
template <int K, int N>
__global__ void __launch_bounds__(256) issue(float *out, int iters, float bb, float cc, long long *cyc,
                                             const __grid_constant__ Rows CR) {
  __shared__ float sm[256];
  sm[threadIdx.x] = bb;
  __syncthreads();
  u64 a[4];
  float2 t;
#pragma unroll
  for (int k = 0; k < 4; ++k) { t = make_float2((float)(threadIdx.x + k), (float)k); a[k] = *reinterpret_cast<u64 *>(&t); }
  t = make_float2(bb, bb); u64 b = *reinterpret_cast<u64 *>(&t);
  t = make_float2(cc, cc); u64 c = *reinterpret_cast<u64 *>(&t);
  float f0 = bb, f1 = cc, f2 = bb + cc, f3 = 1.f;
  int i0 = threadIdx.x, i1 = 3, i2 = 5, i3 = 7;
  unsigned pred = 0;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    const int row0 = (it & 7) * 4;
#pragma unroll
    for (int s = 0; s < 7; ++s) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (K == LDCU_FEED && N > 0 && (s * 4 + k) % (28 / N) == 0) {
          // a row scalar pair through the constant bank into uniform registers, used as the multiplier
          const float4 S = CR.row[row0 * 7 + (s * 4 + k)];
          float2 m = make_float2(S.x, S.x);  // .F32 broadcast operand, like the filter loop
          b = *reinterpret_cast<u64 *>(&m);
        }
        ffma2(a[k], b, c);
        const int e = s * 4 + k;  // slot after the e-th FFMA2
        if (N > 0 && K != LDCU_FEED && e % (28 / N) == 0) {
          // operands change with every FFMA2 (the oldest chain's result), so nothing is loop-invariant
          const unsigned xl = (unsigned)a[(k + 1) & 3];
          const float xf = __uint_as_float(xl);
          if (K == FMNMX) asm volatile("max.f32 %0, %0, %1;" : "+f"(f0) : "f"(xf));
          if (K == FSETP) asm volatile("{ .reg .pred p; setp.ge.f32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(pred) : "f"(xf), "f"(f1));
          if (K == ISETP) asm volatile("{ .reg .pred p; setp.ge.s32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(pred) : "r"(xl), "r"(i1));
          if (K == IADD) asm volatile("add.s32 %0, %0, %1;" : "+r"(i0) : "r"(xl));
          if (K == LOP) asm volatile("xor.b32 %0, %0, %1;" : "+r"(i0) : "r"(xl));
          if (K == IMAD) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i0) : "r"(i1), "r"(xl));
          if (K == VOTE) asm volatile("{ .reg .pred p, q; setp.ge.s32 p, %1, %2; vote.sync.any.pred q, p, 0xffffffff; @q add.u32 %0, %0, 1; }" : "+r"(pred) : "r"(xl), "r"(i1));
          if (K == FMUL) asm volatile("mul.f32 %0, %0, %1;" : "+f"(f0) : "f"(xf));
          if (K == FADD2) asm volatile("add.f32 %0, %0, %1;" : "+f"(f0) : "f"(xf));
          // a never-taken branch around a loop (cannot be if-converted)
          if (K == UBRA) asm volatile("{ .reg .pred p, q; setp.ne.u32 p, %1, 0x7fffffff; @p bra SKIP%=; LOOP%=: add.u32 %0, %0, 3; setp.lt.u32 q, %0, 1000; @q bra LOOP%=; SKIP%=: }" : "+r"(pred) : "r"(xl));
        }
      }
    }
  }
  long long t1 = clock64();
  float s = (float)pred + f0 + f1 + f2 + (float)(i0 + i1 + i2 + i3);
#pragma unroll
  for (int k = 0; k < 4; ++k) { float2 v = *reinterpret_cast<float2 *>(&a[k]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// ---- the filter loop's decision on one group of 4 rows x 2 rays, next to its 28 FFMA2:
// SHAPE 0: max-reduction (2 FMNMX3 + 2 FMNMX) + 2 FSETP + vote + branch (the production shape);
// SHAPE 1: 8 FSETP chained with .OR + vote + branch;  SHAPE 2: no decision at all.
template <int SHAPE>
__global__ void __launch_bounds__(256) decide(float *out, int iters, float bb, float cc, float thr_x, float thr_y) {
  u64 a[4];
  float2 t;
#pragma unroll
  for (int k = 0; k < 4; ++k) { t = make_float2((float)(threadIdx.x + k), (float)k); a[k] = *reinterpret_cast<u64 *>(&t); }
  t = make_float2(bb, bb); u64 b = *reinterpret_cast<u64 *>(&t);
  t = make_float2(cc, cc); u64 c = *reinterpret_cast<u64 *>(&t);
  float sink = 0.f;
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    float2 e[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) e[k] = *reinterpret_cast<float2 *>(&a[k]);  // the previous group's results
#pragma unroll
    for (int s = 0; s < 7; ++s)
#pragma unroll
      for (int k = 0; k < 4; ++k) ffma2(a[k], b, c);
    bool any = false;
    if (SHAPE == 0) {
      const float mx = fmaxf(fmaxf(e[0].x, e[1].x), fmaxf(e[2].x, e[3].x));
      const float my = fmaxf(fmaxf(e[0].y, e[1].y), fmaxf(e[2].y, e[3].y));
      any = (mx >= thr_x) | (my >= thr_y);
    } else if (SHAPE == 1) {
#pragma unroll
      for (int k = 0; k < 4; ++k) any = any | (e[k].x >= thr_x) | (e[k].y >= thr_y);
    }
    if (SHAPE != 2 && __any_sync(0xffffffffu, any)) sink += e[0].x;  // the rare path
  }
  float sacc = sink;
#pragma unroll
  for (int k = 0; k < 4; ++k) { float2 v = *reinterpret_cast<float2 *>(&a[k]); sacc += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = sacc;
}

template <int SHAPE>
void run_decide(const char *name, int sms, int blocks_per_sm, double sm_hz) {
  int blocks = sms * blocks_per_sm, iters = 20000;
  float *out;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    decide<SHAPE><<<blocks, 256>>>(out, iters, 0.999f, 0.001f, 1e30f, 1e30f);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  const double warps_per_sched = blocks_per_sm * 8 / 4.0;
  printf("%-44s %d warps/sched  %6.1f cycles per group of 28 FFMA2\n", name, (int)warps_per_sched,
         best * 1e-3 * sm_hz / iters / warps_per_sched);
  cudaFree(out);
}

template <int K, int N>
void run(const char *name, int sms, int blocks_per_sm, double sm_hz) {
  int blocks = sms * blocks_per_sm, iters = 20000;
  float *out; long long *cyc;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  CK(cudaMalloc(&cyc, blocks * 8));
  static Rows h;
  for (int k = 0; k < 256; ++k) h.row[k] = make_float4(0.999f, 0.999f, 0.5f, 0.25f);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    issue<K, N><<<blocks, 256>>>(out, iters, 0.999f, 0.001f, cyc, h);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  // one resident wave: every scheduler runs warps_per_sched warps for the whole kernel
  const double warps_per_sched = blocks_per_sm * 8 / 4.0;
  const double cyc_per_iter = best * 1e-3 * sm_hz / iters / warps_per_sched;
  printf("%-22s +%2d per 28 FFMA2  %d warps/sched  %6.1f cycles/iter/warp  (+%.2f per extra instruction)\n", name, N,
         (int)warps_per_sched, cyc_per_iter, N ? (cyc_per_iter - 56.0) / N : 0.0);
  cudaFree(out); cudaFree(cyc);
}

#define SWEEP(K, name) run<K, 4>(name, sms, bps, hz); run<K, 14>(name, sms, bps, hz); run<K, 28>(name, sms, bps, hz);

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  const double hz = p.clockRate * 1e3;  // the maximum SM clock; a B200 holds it through this run
  printf("%s, %d SMs, %.0f MHz\n", p.name, sms, hz / 1e6);
  for (int bps : {1, 3}) {
    run<NONE, 0>("FFMA2 only", sms, bps, hz);
    SWEEP(FMNMX, "FMNMX");
    SWEEP(FSETP, "FSETP(+@p IADD)");
    SWEEP(ISETP, "ISETP(+@p IADD)");
    SWEEP(IADD, "IADD3");
    SWEEP(LOP, "LOP3");
    SWEEP(IMAD, "IMAD");
    SWEEP(FMUL, "FMUL");
    SWEEP(FADD2, "FADD");
    SWEEP(VOTE, "ISETP+VOTE(+@q IADD)");
    SWEEP(UBRA, "ISETP+BRA");
    SWEEP(LDCU_FEED, "LDCU.64 feed");
    run_decide<2>("group: no decision", sms, bps, hz);
    run_decide<0>("group: max-reduce + 2 FSETP + vote + branch", sms, bps, hz);
    run_decide<1>("group: 8 FSETP.OR + vote + branch", sms, bps, hz);
  }
  return 0;
}
