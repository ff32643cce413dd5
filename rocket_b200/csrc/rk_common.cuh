// Shared definitions for librocket_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/rocket_b200.h"

#ifndef __CUDA_ARCH__
#define RK_HOST 1
#endif

// ---- error plumbing ---------------------------------------------------------
void rk_set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_rk_launches;

#define RK_REQUIRE(cond, code, ...)    \
  do {                                 \
    if (!(cond)) {                     \
      rk_set_error(__VA_ARGS__);       \
      return (code);                   \
    }                                  \
  } while (0)

#define RK_CUDA(expr)                                                        \
  do {                                                                       \
    cudaError_t _e = (expr);                                                 \
    if (_e != cudaSuccess) {                                                 \
      rk_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
                   __FILE__, __LINE__);                                      \
      return (int)_e;                                                        \
    }                                                                        \
  } while (0)

// ---- plan blob ---------------------------------------------------------------
// Layout: [rk_plan_header][descs n][prefix_step n+1][prefix_norm n+1], each section
// 16-byte aligned.  Offsets are relative to the blob start so the same bytes are
// valid on host and device.
#define RK_PLAN_MAGIC 0x524b5031u /* "RKP1" */

struct rk_plan_header {
  uint32_t magic;
  int32_t n_tensors;
  int32_t chunk_step;     // elements per step chunk (RK_STEP_CHUNK)
  int32_t chunk_norm;     // elements per norm chunk (power of two >= RK_NORM_CHUNK_MIN)
  int32_t n_chunks_step;
  int32_t n_chunks_norm;
  uint32_t off_descs;
  uint32_t off_prefix_step;
  uint32_t off_prefix_norm;
  uint32_t plan_bytes;
  // scratch layout (bytes from scratch base)
  uint32_t off_partials;  // double[n_chunks_norm]
  uint32_t off_tnorms;    // float[n_tensors]
  uint32_t off_total;     // float[1]
  uint32_t off_ticket;    // uint32[1]  (must be zero before the first launch)
  uint32_t scratch_bytes;
  uint32_t _pad;
  int64_t total_numel;
};

// Tensors whose four pointers are 16-byte aligned take the 128-bit path.
#define RK_TENSOR_VEC4 (1 << 8)
// ... 32-byte aligned: the 256-bit path.
#define RK_TENSOR_VEC8 (1 << 9)

// How many prefix entries a CTA stages in shared memory; longer lists are searched
// in global memory (L1/L2 resident).
#define RK_SMEM_PREFIX_CAP 4096

static inline size_t rk_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- device helpers -----------------------------------------------------------
#ifdef __CUDACC__

// Stage the chunk-prefix table (tensor-list metadata) into shared memory.
__device__ __forceinline__ const int32_t* rk_stage_prefix(const int32_t* __restrict__ gprefix,
                                                          int n_entries, int32_t* smem) {
  if (n_entries > RK_SMEM_PREFIX_CAP) return gprefix;
  for (int i = threadIdx.x; i < n_entries; i += blockDim.x) smem[i] = __ldg(gprefix + i);
  __syncthreads();
  return smem;
}

// Largest t with prefix[t] <= c   (prefix[0] = 0, prefix[n] = n_chunks, c < n_chunks).
__device__ __forceinline__ int rk_find_tensor(const int32_t* prefix, int n_tensors, int c) {
  int lo = 0, hi = n_tensors;  // invariant: prefix[lo] <= c < prefix[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (prefix[mid] <= c) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ float rk_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double rk_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// clip_coef_clamped of $TORCH/nn/utils/clip_grad.py:165-169, in fp32 like torch:
// reciprocal(total + 1e-6) * max_norm, clamp(max=1) (NaN propagates).
__device__ __forceinline__ float rk_clip_coef(float total_norm, float max_norm) {
  float c = __fmul_rn(__frcp_rn(__fadd_rn(total_norm, 1e-6f)), max_norm);
  return c > 1.0f ? 1.0f : c;
}

#endif  // __CUDACC__

// Grid for a persistent kernel: CTAs per SM (by occupancy) x SM count, capped by work.
int rk_persistent_grid(const void* kernel, int block, size_t dyn_smem, int n_work, int* grid_out);
