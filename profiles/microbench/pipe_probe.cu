// pipe_probe.cu -- design microbenchmarks for the rollout kernel on sm_100a (B200):
// issue/pipe throughput of FFMA (uniform operand), FFMA (3 registers), FFMA2 (packed f32x2,
// Blackwell), MUFU.EX2/RCP, and an NN-like mix.  Build: nvcc -gencode arch=compute_100a,code=sm_100a
// -O3 -o pipe_probe pipe_probe.cu ; run on the GPU box.  Prints lane-op rates per clk per SM.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack(float lo, float hi){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(lo),"f"(hi)); return r;}
__device__ __forceinline__ void unpack(u64 v, float&lo, float&hi){ asm("mov.b64 {%0,%1}, %2;":"=f"(lo),"=f"(hi):"l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ float ex2(float x){float y; asm("ex2.approx.ftz.f32 %0, %1;":"=f"(y):"f"(x)); return y;}
__device__ __forceinline__ float rcp(float x){float y; asm("rcp.approx.ftz.f32 %0, %1;":"=f"(y):"f"(x)); return y;}

#define N 16
template <int MODE>
__global__ void __launch_bounds__(256) probe(float* out, int iters, float a, float b) {
  float acc[N]; u64 acc2[N];
#pragma unroll
  for (int i = 0; i < N; ++i) { acc[i] = (threadIdx.x + i) * 1e-3f; acc2[i] = pack(acc[i], acc[i] + 1.f); }
  float r0 = a * threadIdx.x, r1 = b + threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (MODE == 0) {            // FFMA, uniform multiplier + register addend
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] = fmaf(acc[i], a, b);
      } else if (MODE == 1) {     // FFMA, three distinct registers
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] = fmaf(acc[i], acc[(i + 5) % N], acc[(i + 11) % N]);
      } else if (MODE == 2) {     // FFMA2, broadcast uniform multiplier
        const u64 aa = pack(a, a), bb = pack(b, b);
#pragma unroll
        for (int i = 0; i < N; ++i) acc2[i] = ffma2(acc2[i], aa, bb);
      } else if (MODE == 3) {     // FFMA2, three packed registers
#pragma unroll
        for (int i = 0; i < N; ++i) acc2[i] = ffma2(acc2[i], acc2[(i + 5) % N], acc2[(i + 11) % N]);
      } else if (MODE == 4) {     // MUFU.EX2
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] = ex2(acc[i]);
      } else if (MODE == 5) {     // NN-like mix, scalar: 11 FFMA + EX2 + FADD + RCP per unit
#pragma unroll
        for (int i = 0; i < N; ++i) {
          float h = fmaf(acc[i], a, b); h = fmaf(r0, a, h); h = fmaf(r1, b, h); h = fmaf(r0, b, h); h = fmaf(r1, a, h);
          float s = rcp(1.0f + ex2(h));
          acc[i] = fmaf(s, a, acc[i]); acc[(i + 1) % N] = fmaf(s, b, acc[(i + 1) % N]);
          acc[(i + 2) % N] = fmaf(s, a, acc[(i + 2) % N]); acc[(i + 3) % N] = fmaf(s, b, acc[(i + 3) % N]);
          acc[(i + 4) % N] = fmaf(s, a, acc[(i + 4) % N]); acc[(i + 5) % N] = fmaf(s, b, acc[(i + 5) % N]);
        }
      } else if (MODE == 6) {     // NN-like mix, packed: 11 FFMA2 + 2 EX2 + FADD2 + 2 RCP per unit pair
        const u64 aa = pack(a, a), bb = pack(b, b), one = pack(1.f, 1.f), R0 = pack(r0, r1), R1 = pack(r1, r0);
#pragma unroll
        for (int i = 0; i < N; ++i) {
          u64 h = ffma2(acc2[i], aa, bb); h = ffma2(R0, aa, h); h = ffma2(R1, bb, h); h = ffma2(R0, bb, h); h = ffma2(R1, aa, h);
          float h0, h1; unpack(h, h0, h1);
          u64 d = ffma2(pack(ex2(h0), ex2(h1)), one, one);
          float d0, d1; unpack(d, d0, d1);
          const u64 s = pack(rcp(d0), rcp(d1));
          acc2[i] = ffma2(s, aa, acc2[i]); acc2[(i + 1) % N] = ffma2(s, bb, acc2[(i + 1) % N]);
          acc2[(i + 2) % N] = ffma2(s, aa, acc2[(i + 2) % N]); acc2[(i + 3) % N] = ffma2(s, bb, acc2[(i + 3) % N]);
          acc2[(i + 4) % N] = ffma2(s, aa, acc2[(i + 4) % N]); acc2[(i + 5) % N] = ffma2(s, bb, acc2[(i + 5) % N]);
        }
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < N; ++i) { float lo, hi; unpack(acc2[i], lo, hi); s += acc[i] + lo + hi; }
  if (s == 123.456f) out[0] = s;
}

template <int MODE> double run(int iters, int warps_per_smsp, int sms) {
  float* d; cudaMalloc(&d, 256);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int threads = 256, blocks = sms * warps_per_smsp * 4 * 32 / threads;
  probe<MODE><<<blocks, threads>>>(d, iters, 0.999f, 1e-4f);
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); probe<MODE><<<blocks, threads>>>(d, iters, 0.999f, 1e-4f); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float t; cudaEventElapsedTime(&t, e0, e1); if (t < best) best = t;
  }
  cudaFree(d);
  return (double)blocks * threads * iters * 4.0 * N / (best * 1e-3);   // inner-loop "units" per second
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const double clk = clk_khz * 1e3, sms = p.multiProcessorCount;
  printf("%s  SMs %d  clock %.0f MHz (max)\n", p.name, p.multiProcessorCount, clk / 1e6);
  const int it = 2048;
  for (int w = 2; w <= 8; w *= 2) {
    printf("warps/SMSP %d\n", w);
    double u;
    u = run<0>(it, w, sms); printf("  FFMA  (R,UR,R)      %7.1f FMA-lanes/clk/SM   %6.2f TFLOP/s\n", u / clk / sms, 2 * u / 1e12);
    u = run<1>(it, w, sms); printf("  FFMA  (R,R,R)       %7.1f FMA-lanes/clk/SM   %6.2f TFLOP/s\n", u / clk / sms, 2 * u / 1e12);
    u = run<2>(it, w, sms); printf("  FFMA2 (R2,UR.bc,R2) %7.1f FMA-lanes/clk/SM   %6.2f TFLOP/s\n", 2 * u / clk / sms, 4 * u / 1e12);
    u = run<3>(it, w, sms); printf("  FFMA2 (R2,R2,R2)    %7.1f FMA-lanes/clk/SM   %6.2f TFLOP/s\n", 2 * u / clk / sms, 4 * u / 1e12);
    u = run<4>(it / 4, w, sms); printf("  MUFU.EX2            %7.1f lanes/clk/SM\n", u / clk / sms);
    u = run<5>(it / 8, w, sms); printf("  NN mix scalar       %7.2f unit-lanes/clk/SM  (11 FMA + 2 MUFU + 1 FADD each) -> %6.2f TFLOP/s FMA\n", u / clk / sms, 22 * u / 1e12);
    u = run<6>(it / 8, w, sms); printf("  NN mix packed       %7.2f unit-lanes/clk/SM  (per trajectory)               -> %6.2f TFLOP/s FMA\n", 2 * u / clk / sms, 44 * u / 1e12);
  }
  return 0;
}
