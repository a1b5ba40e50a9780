// Shared launch bookkeeping for every extern "C" entry point of libbiscotti_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace bsc {

// number of kernel launches issued through this library (bench.py reports it as gpu_launches)
extern unsigned long long g_launch_count;

inline int finish_launch(const char* what, int nlaunch = 1) {
  g_launch_count += (unsigned long long)nlaunch;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    fprintf(stderr, "[biscotti_b200] %s: %s\n", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

#define BSC_CHECK(call)                                                                   \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess) {                                                              \
      fprintf(stderr, "[biscotti_b200] %s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return (int)_e;                                                                     \
    }                                                                                     \
  } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum, result broadcast to all threads; `red` must hold >= 33 floats
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (warp == 0) {
    float t = (lane < (int)((blockDim.x + 31) >> 5)) ? red[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

}  // namespace bsc
