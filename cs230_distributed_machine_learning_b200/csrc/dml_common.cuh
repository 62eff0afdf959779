// Shared helpers for the sweep kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#define DML_OK 0
#define DML_ERR_ARG 1
#define DML_ERR_CUDA 2
#define DML_ERR_NCCL 3
#define DML_ERR_LIMIT 4

namespace dml {

constexpr int kMaxCor = 10;        // L-BFGS-B memory (scipy default maxcor)
constexpr int kMaxLs = 50;         // sklearn passes maxls = 50 (_logistic.py:589)
constexpr double kEps = 2.220446049250313e-16;
constexpr double kFactr = 64.0;    // sklearn passes ftol = 64*eps
constexpr double kStpMax = 1e10;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum / max of doubles, fixed order => deterministic. All threads
// must call; result is returned to every thread. `red` holds >= 33 doubles.
template <bool IS_MAX>
__device__ __forceinline__ double block_reduce(double v, double* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  __syncthreads();  // protect `red` from the previous call's readers
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    double t = lane < nw ? red[lane] : (IS_MAX ? -INFINITY : 0.0);
    t = IS_MAX ? warp_max(t) : warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

}  // namespace dml
