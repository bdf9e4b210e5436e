// airc_peak.cuh -- FP32 FFMA throughput probe.  MEASURED_PEAKS.json carries HBM and bf16-tensor
// peaks only; the rollout kernel is bound by the CUDA-core FMA pipe, so bench.py measures that
// roofline denominator on the same device, in the same run, under the same clocks.
#pragma once
#include <cuda_runtime.h>

namespace airc {

// 16 independent accumulators per thread, 2 FMA each per inner iteration: long enough dependent
// distance (16 >> FFMA latency 4) for 1 FFMA/clk/lane with >= 4 resident warps per SMSP.
__global__ void __launch_bounds__(256) ffma_peak_kernel(float* out, int iters, float a, float b) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = (float)(threadIdx.x + i) * 1e-3f;
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  if (s == 123.456f) out[0] = s;   // never true; keeps the chain alive
}

}  // namespace airc
