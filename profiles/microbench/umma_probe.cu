// umma_probe.cu -- feasibility probe for carrying the surrogate's two layers on tcgen05 (TF32, A operand in TMEM):
//   1. correctness of the hand-built descriptors: D[128 x N] = A[128 x K] . B[N x K]^T, A written by the threads
//      with tcgen05.st (one row per thread), B in shared memory (K-major, no swizzle), D read with tcgen05.ld;
//   2. accuracy of the 3-term TF32 split (a_hi b_hi + a_lo b_hi + a_hi b_lo) against float64;
//   3. round-trip latency st -> fence -> barrier -> mma -> commit -> mbarrier wait -> ld, in cycles.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_probe umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#define DEVI __device__ __forceinline__
DEVI uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 bytes, contiguous (128 B);
// SBO = byte stride between 8-row groups, LBO = byte stride between 16-byte K chunks.
DEVI uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
  return d;                                      // layout type 0 = no swizzle, base offset 0
}
// instruction descriptor: D F32, A/B TF32, both K-major
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
DEVI void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
DEVI void commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
DEVI bool bar_wait(uint32_t bar, uint32_t parity) {   // bounded: returns false on time-out instead of hanging the box
  for (int spin = 0; spin < (1 << 22); ++spin) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}
DEVI void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DEVI void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
DEVI void st8(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                  "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                  "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])) : "memory");
}
DEVI void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DEVI void ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

constexpr int N = 16, K = 8;     // one MMA: M 128 x N 16 x K 8
// Bsm layout (floats): chunk c = k/4 at c*(N/8)*32 floats; 8-row group g = n/8 at g*32; row r = n%8 at r*4; k%4
DEVI int b_index(int n, int k) { return (k >> 2) * (N / 8) * 32 + (n >> 3) * 32 + (n & 7) * 4 + (k & 3); }

// mode 0: plain (A, B as given: TF32-exact inputs expected); mode 1: 3-term split of arbitrary FP32 inputs
__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, int mode, int reps, long long* cyc,
                                             int* status) {
  __shared__ __align__(128) float Bs[2][N * K];       // [0] hi, [1] lo
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" :: "r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < N * K; i += 128) {
    const int n = i / K, k = i % K;
    const float b = B[n * K + k];
    const float bh = __uint_as_float(__float_as_uint(b) & 0xFFFFE000u);
    Bs[0][b_index(n, k)] = mode ? bh : b;
    Bs[1][b_index(n, k)] = b - bh;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the MMA (async proxy)
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = tmem_base_s;
  const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
  const uint32_t colD = 0, colAh = 16, colAl = 24;
  float a[K], ah[K], al[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    a[k] = A[tid * K + k];
    ah[k] = __uint_as_float(__float_as_uint(a[k]) & 0xFFFFE000u);
    al[k] = a[k] - ah[k];
  }
  const uint64_t dBh = make_desc(smem_u32(&Bs[0][0]), (N / 8) * 128, 128);
  const uint64_t dBl = make_desc(smem_u32(&Bs[1][0]), (N / 8) * 128, 128);
  const uint32_t idesc = make_idesc(128, N);
  float d[16];
  uint32_t parity = 0;
  bool ok = true;
  long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    st8(lane_base + colAh, mode ? ah : a);
    if (mode) st8(lane_base + colAl, al);
    wait_st();
    fence_before();
    __syncthreads();
    if (tid == 0) {
      fence_after();
      mma_ts(tbase + colD, tbase + colAh, dBh, idesc, 0);
      if (mode) {
        mma_ts(tbase + colD, tbase + colAl, dBh, idesc, 1);
        mma_ts(tbase + colD, tbase + colAh, dBl, idesc, 1);
      }
      commit(smem_u32(&mbar));
    }
    ok = bar_wait(smem_u32(&mbar), parity) && ok;
    parity ^= 1;
    fence_after();
    ld16(lane_base + colD, d);
    if (r + 1 < reps) a[0] += d[0] * 1e-30f;         // dependency chain for the latency figure
  }
  long long t1 = clock64();
  if (tid == 0) { cyc[0] = (t1 - t0) / reps; status[0] = ok ? 0 : 1; }
#pragma unroll
  for (int n = 0; n < N; ++n) D[tid * N + n] = d[n];
  fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" :: "r"(tbase) : "memory");
}

int main() {
  float hA[128 * K], hB[N * K], hD[128 * N];
  float *dA, *dB, *dD; long long* dc; int* ds;
  cudaMalloc(&dA, sizeof(hA)); cudaMalloc(&dB, sizeof(hB)); cudaMalloc(&dD, sizeof(hD)); cudaMalloc(&dc, 8); cudaMalloc(&ds, 4);
  for (int mode = 0; mode < 2; ++mode) {
    srand(1 + mode);
    for (auto& v : hA) v = mode ? (float)(rand() / (double)RAND_MAX * 2 - 1) : (float)((rand() % 33) - 16) / 8.0f;
    for (auto& v : hB) v = mode ? (float)(rand() / (double)RAND_MAX * 2 - 1) : (float)((rand() % 33) - 16) / 16.0f;
    cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, sizeof(hB), cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xFF, sizeof(hD));
    probe<<<1, 128>>>(dA, dB, dD, mode, 1, dc, ds);
    cudaError_t e = cudaDeviceSynchronize();
    int st = -1; cudaMemcpy(&st, ds, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(hD, dD, sizeof(hD), cudaMemcpyDeviceToHost);
    double maxerr = 0, maxref = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < N; ++n) {
        double ref = 0;
        for (int k = 0; k < K; ++k) ref += (double)hA[m * K + k] * (double)hB[n * K + k];
        maxerr = fmax(maxerr, fabs(ref - (double)hD[m * N + n])); maxref = fmax(maxref, fabs(ref));
      }
    printf("mode %d (%s): cuda %s, barrier %s, max |D - ref| = %.3e (max |ref| %.3f)  D[0][0..3] = %g %g %g %g\n", mode,
           mode ? "3-term TF32 split, random FP32 inputs" : "TF32-exact inputs", cudaGetErrorString(e), st == 0 ? "ok" : "TIMEOUT",
           maxerr, maxref, hD[0], hD[1], hD[2], hD[3]);
    if (e != cudaSuccess) return 1;
  }
  for (int mode = 0; mode < 2; ++mode) {
    probe<<<1, 128>>>(dA, dB, dD, mode, 2000, dc, ds);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    printf("round trip (st -> barrier -> %d mma -> commit -> wait -> ld), one CTA alone: %lld cycles\n", mode ? 3 : 1, c);
  }
  return 0;
}
