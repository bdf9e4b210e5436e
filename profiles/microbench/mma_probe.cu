// mma_probe.cu -- can the legacy tensor path (mma.sync, SASS HMMA) carry the surrogate's two tiny layers?
// Measures on sm_100a: mma.sync.m16n8k8 TF32 issue rate per SM (independent accumulators), alone and interleaved
// with FFMA2 / MUFU work, at 4 / 8 / 12 warps per SM (the sweep kernel runs 12).  The question is whether 96 MMAs
// per warp per RK4 stage (3xTF32 split of a 64x8x32 and a 64x32x8 product) cost less FMA-pipe / issue time than
// the 352 FFMA2 they would replace (704 FMA-pipe cycles per warp per stage).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe mma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ void mma_tf32(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

#define NACC 8
// MODE 0: MMA only; 1: MMA + 2 FFMA2 per MMA; 2: MMA + 4 FFMA2 per MMA; 3: FFMA2 only (4 per slot); 4: MMA + 1 MUFU per MMA
template <int MODE>
__global__ void probe(float* out, int iters, float s) {
  float d[NACC][4];
  unsigned a[4], b[2];
  u64 acc[8];
  float m = s * threadIdx.x;
#pragma unroll
  for (int i = 0; i < NACC; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) a[j] = __float_as_uint(s * (threadIdx.x + j));
  b[0] = __float_as_uint(s); b[1] = __float_as_uint(2 * s);
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = pack(s * i, s + i);
  const u64 aa = pack(s, s), bb = pack(1e-3f, 1e-3f);
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (MODE != 3) mma_tf32(d[i], a, b);
      if (MODE == 1) { acc[i] = ffma2(acc[i], aa, bb); acc[(i + 4) & 7] = ffma2(acc[(i + 4) & 7], aa, bb); }
      if (MODE == 2 || MODE == 3) {
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[(i + 2 * k) & 7] = ffma2(acc[(i + 2 * k) & 7], aa, bb);
      }
      if (MODE == 4) m = ex2(m);
    }
  }
  float r = m;
#pragma unroll
  for (int i = 0; i < NACC; ++i) r += d[i][0] + d[i][1] + d[i][2] + d[i][3];
#pragma unroll
  for (int i = 0; i < 8; ++i) r += (float)(acc[i] & 0xffff);
  if (r == 123.456f) out[0] = r;
}

template <int MODE>
static void run(const char* name, int warps_per_sm, int sms, double mhz) {
  float* out; cudaMalloc(&out, 4);
  const int iters = 20000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<MODE><<<sms, warps_per_sm * 32>>>(out, 100, 1e-3f);
  cudaEventRecord(e0);
  probe<MODE><<<sms, warps_per_sm * 32>>>(out, iters, 1e-3f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double clk = ms * 1e-3 * mhz * 1e6;                     // cycles
  const double slots = (double)iters * NACC * warps_per_sm;     // per SM
  printf("%-34s warps/SM %2d: %7.3f ms  %.3f slots/clk/SM  (%.2f clk per slot per SMSP)\n", name, warps_per_sm, ms,
         slots / clk, clk / (slots / 4));
  cudaFree(out);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const double mhz = p.clockRate / 1000.0;
  printf("%s, %d SMs, %.0f MHz nominal; one slot = 1 mma.m16n8k8.tf32 (+ the listed extra work)\n", p.name, p.multiProcessorCount, mhz);
  for (int w : {4, 8, 12, 16}) {
    run<0>("mma only", w, p.multiProcessorCount, mhz);
    run<1>("mma + 2 FFMA2", w, p.multiProcessorCount, mhz);
    run<2>("mma + 4 FFMA2", w, p.multiProcessorCount, mhz);
    run<3>("4 FFMA2 only", w, p.multiProcessorCount, mhz);
    run<4>("mma + 1 MUFU.EX2", w, p.multiProcessorCount, mhz);
  }
  return 0;
}
