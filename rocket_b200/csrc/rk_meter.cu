// K4: Meter reductions on the device (no per-step host sync).
//   rk_meter_argmax_count : counters[0] += #(argmax(logits[r,:]) == labels[r]), counters[1] += rows
//   rk_scalar_accumulate  : *acc += (double)*value / divisor
// Integer atomics only, so the counts are exact and independent of scheduling.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "rk_common.cuh"

#define METER_THREADS 256
#define METER_WARPS (METER_THREADS / 32)

template <typename T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_f32<__half>(__half x) { return __half2float(x); }

// torch.argmax ordering: NaN is the maximum; among equals the lowest index wins.
__device__ __forceinline__ bool better(float a, int64_t ia, float b, int64_t ib) {
  const bool an = a != a, bn = b != b;
  if (an || bn) return an && (!bn || ia < ib);
  return a > b || (a == b && ia < ib);
}

template <typename T>
__global__ void __launch_bounds__(METER_THREADS)
rk_argmax_count_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                       int64_t n_rows, int64_t n_cols, int64_t row_stride,
                       unsigned long long* __restrict__ counters) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int hits = 0;
  for (int64_t r = (int64_t)blockIdx.x * METER_WARPS + warp; r < n_rows;
       r += (int64_t)gridDim.x * METER_WARPS) {
    const T* __restrict__ row = logits + r * row_stride;
    float best = -INFINITY;
    int64_t bi = INT64_MAX;  // any real index beats the sentinel on equality
    for (int64_t j = lane; j < n_cols; j += 32) {
      const float x = to_f32<T>(row[j]);
      if (better(x, j, best, bi)) { best = x; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (better(ob, oi, best, bi)) { best = ob; bi = oi; }
    }
    if (lane == 0 && n_cols > 0 && bi == labels[r]) ++hits;
  }
  __shared__ int s_hits[METER_WARPS];
  if (lane == 0) s_hits[warp] = hits;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
#pragma unroll
    for (int w = 0; w < METER_WARPS; ++w) tot += s_hits[w];
    if (tot) atomicAdd(counters + 0, (unsigned long long)tot);
    if (blockIdx.x == 0) atomicAdd(counters + 1, (unsigned long long)n_rows);
  }
}

extern "C" int rk_meter_argmax_count(const void* logits, int dtype, const int64_t* labels,
                                     int64_t n_rows, int64_t n_cols, int64_t row_stride,
                                     long long* counters_dev, void* stream) {
  RK_REQUIRE(counters_dev, RK_ERR_INVALID, "rk_meter_argmax_count: null counters");
  RK_REQUIRE(n_rows >= 0 && n_cols >= 0, RK_ERR_INVALID,
             "rk_meter_argmax_count: negative shape (%lld x %lld)", (long long)n_rows,
             (long long)n_cols);
  if (n_rows == 0) return 0;
  RK_REQUIRE(logits && labels, RK_ERR_INVALID, "rk_meter_argmax_count: null logits/labels");
  RK_REQUIRE(row_stride >= n_cols, RK_ERR_INVALID,
             "rk_meter_argmax_count: row_stride %lld < n_cols %lld", (long long)row_stride,
             (long long)n_cols);
  int64_t blocks = (n_rows + METER_WARPS - 1) / METER_WARPS;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cudaStream_t s = (cudaStream_t)stream;
  unsigned long long* ctr = (unsigned long long*)counters_dev;
  switch (dtype) {
    case 0:
      rk_argmax_count_kernel<float><<<(int)blocks, METER_THREADS, 0, s>>>(
          (const float*)logits, labels, n_rows, n_cols, row_stride, ctr);
      break;
    case 1:
      rk_argmax_count_kernel<__nv_bfloat16><<<(int)blocks, METER_THREADS, 0, s>>>(
          (const __nv_bfloat16*)logits, labels, n_rows, n_cols, row_stride, ctr);
      break;
    case 2:
      rk_argmax_count_kernel<__half><<<(int)blocks, METER_THREADS, 0, s>>>(
          (const __half*)logits, labels, n_rows, n_cols, row_stride, ctr);
      break;
    default:
      RK_REQUIRE(false, RK_ERR_UNSUPPORTED, "rk_meter_argmax_count: dtype %d (0=f32,1=bf16,2=f16)",
                 dtype);
  }
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
}

__global__ void rk_scalar_accumulate_kernel(const float* __restrict__ value, double divisor,
                                            double* __restrict__ acc) {
  *acc = *acc + (double)(*value) / divisor;
}

extern "C" int rk_scalar_accumulate(const float* value_dev, double divisor, double* acc_dev,
                                    void* stream) {
  RK_REQUIRE(value_dev && acc_dev, RK_ERR_INVALID, "rk_scalar_accumulate: null pointer");
  RK_REQUIRE(divisor != 0.0, RK_ERR_INVALID, "rk_scalar_accumulate: divisor is zero");
  rk_scalar_accumulate_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(value_dev, divisor, acc_dev);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
}
