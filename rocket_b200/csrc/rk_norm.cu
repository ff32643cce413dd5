// K1: deterministic multi-tensor gradient L2 norm (HBM-bound: 4 B read per element).
//
// Level 1 (rk_l2norm_kernel): every norm chunk (plan.chunk_norm elements of one tensor) ->
//          one double partial, written to scratch in chunk order.  8 CTAs per SM.
// Level 2 (rk_l2norm_finalize_kernel): ONE CTA of 1024 threads reduces, in a fixed order,
//          partials -> per-tensor norms -> total norm.  It is its own launch -- chained behind
//          level 1 (or behind the fused exchange kernel, which writes the same partials) as a
//          programmatic dependent, so its launch latency and set-up hide behind the producer's
//          tail -- because it wants what level 1 cannot spare inside an 8-CTA/SM kernel:
//          registers for eight loads in flight per thread and 32 warps.  (Inlined into level 1's
//          last CTA it was a chain of dependent round trips with 4 loads in flight and 8 warps:
//          12 us on a ResNet-18, 125 us on a 1,185-tensor list; profiles/r2_norm_notes.md.)
// No float atomics anywhere, so the result does not depend on CTA scheduling and is
// bit-identical on every rank that holds the same gradients.
#include <mutex>

#include "rk_common.cuh"

#define NORM_THREADS 256
#define NORM_WARPS (NORM_THREADS / 32)
#define NORM_UNROLL 4
#define FIN_THREADS 1024
#define FIN_WARPS (FIN_THREADS / 32)
#define FIN_PREFIX_CAP 2048   // prefix entries staged in shared memory (8 KB: the CTA must fit
                              // beside a resident step-kernel CTA, which leaves ~17 KB per SM)
#define FIN_LIST_CAP 1024     // work-list entries (tensors one warp reduces)
#define FIN_BIG_CAP 128       // ... (tensors the whole CTA reduces)
#define NORM_BIG_TENSOR 1024  // tensors with more chunks than this are reduced block-wide
#define NORM_THREAD_TENSOR 8  // tensors with at most this many chunks are summed by one thread

__device__ __forceinline__ float sq_acc(float acc, float x, float s) {
  float y = x * s;
  return fmaf(y, y, acc);
}

// Sum of partials[c0 + first], partials[c0 + first + stride], ... below c1: eight independent
// accumulators (eight loads in flight), combined in a fixed tree.
__device__ __forceinline__ double strided_sum8(const double* __restrict__ partials, int c0, int c1,
                                               int first, int stride) {
  double s[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  int c = c0 + first;
  for (; c + 7 * stride < c1; c += 8 * stride) {
    double x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = __ldcg(partials + c + k * stride);
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] += x[k];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (c < c1) s[k] += __ldcg(partials + c);
    c += stride;
  }
  return ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
}

// ---- level 2: partials -> per-tensor norms -> total norm, by ONE CTA, in a fixed order ----
__global__ void __launch_bounds__(FIN_THREADS, 1)
rk_l2norm_finalize_kernel(const char* __restrict__ plan, char* __restrict__ scratch,
                          float* __restrict__ total_out, float* __restrict__ tnorm_out) {
  __shared__ int32_t s_prefix[FIN_PREFIX_CAP];
  __shared__ double s_red[FIN_WARPS];
  __shared__ int s_mid[FIN_LIST_CAP], s_big[FIN_BIG_CAP];
  __shared__ int s_n_mid, s_n_big;
  if (threadIdx.x == 0) s_n_mid = s_n_big = 0;
  // a step kernel launched as a programmatic dependent may set itself up while this runs
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // the plan is constant: stage the chunk prefix before waiting for the partials' producer
  const int32_t* gp = reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm);
  const int32_t* prefix = gp;
  if (n_tensors + 1 <= FIN_PREFIX_CAP) {
    for (int i = tid; i <= n_tensors; i += FIN_THREADS) s_prefix[i] = __ldg(gp + i);
    prefix = s_prefix;
  }
  __syncthreads();
  asm volatile("griddepcontrol.wait;" ::: "memory");   // partials are the previous kernel's output
  const double* partials = reinterpret_cast<const double*>(scratch + h->off_partials);
  float* tnorms = reinterpret_cast<float*>(scratch + h->off_tnorms);

  // ---- 2a: tensors of at most NORM_THREAD_TENSOR chunks, one THREAD each, in chunk order;
  //          larger ones are collected into two work lists (their order does not matter: each
  //          tensor's norm is computed on its own, in a fixed order) ----
  for (int t = tid; t < n_tensors; t += FIN_THREADS) {
    const int c0 = prefix[t], n = prefix[t + 1] - c0;
    if (n > NORM_THREAD_TENSOR) {
      if (n <= NORM_BIG_TENSOR) {
        const int k = atomicAdd(&s_n_mid, 1);
        if (k < FIN_LIST_CAP) s_mid[k] = t;
      } else {
        const int k = atomicAdd(&s_n_big, 1);
        if (k < FIN_BIG_CAP) s_big[k] = t;
      }
      continue;
    }
    double x[NORM_THREAD_TENSOR];
#pragma unroll
    for (int k = 0; k < NORM_THREAD_TENSOR; ++k) x[k] = k < n ? __ldcg(partials + c0 + k) : 0.0;
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < NORM_THREAD_TENSOR; ++k) s += x[k];
    tnorms[t] = (float)sqrt(s);
  }
  __syncthreads();
  const int n_mid = s_n_mid, n_big = s_n_big;
  // ---- 2b: up to NORM_BIG_TENSOR chunks, one WARP each ---------------------------------------
  if (n_mid <= FIN_LIST_CAP) {
    for (int k = warp; k < n_mid; k += FIN_WARPS) {
      const int t = s_mid[k];
      const double s = rk_warp_sum(strided_sum8(partials, prefix[t], prefix[t + 1], lane, 32));
      if (lane == 0) tnorms[t] = (float)sqrt(s);
    }
  } else {                               // list overflow: scan all tensors
    for (int t = warp; t < n_tensors; t += FIN_WARPS) {
      const int c0 = prefix[t], c1 = prefix[t + 1];
      if (c1 - c0 <= NORM_THREAD_TENSOR || c1 - c0 > NORM_BIG_TENSOR) continue;
      const double s = rk_warp_sum(strided_sum8(partials, c0, c1, lane, 32));
      if (lane == 0) tnorms[t] = (float)sqrt(s);
    }
  }
  // ---- 2c: bigger tensors, the whole CTA each ------------------------------------------------
  const bool big_listed = n_big <= FIN_BIG_CAP;
  const int n_scan = big_listed ? n_big : n_tensors;
  for (int k = 0; k < n_scan; ++k) {
    const int t = big_listed ? s_big[k] : k;
    const int c0 = prefix[t], c1 = prefix[t + 1];
    if (c1 - c0 <= NORM_BIG_TENSOR) continue;
    const double s = rk_warp_sum(strided_sum8(partials, c0, c1, tid, FIN_THREADS));
    if (lane == 0) s_red[warp] = s;
    __syncthreads();
    if (tid == 0) {
      double tot = 0.0;
#pragma unroll
      for (int w = 0; w < FIN_WARPS; ++w) tot += s_red[w];
      tnorms[t] = (float)sqrt(tot);
    }
    __syncthreads();
  }
  __syncthreads();  // tnorms written by this CTA are visible to this CTA
  // ---- 2d: norm of the per-tensor norms (clip_grad.py:106-107) ---------------------------
  double acc = 0.0;
  for (int t = tid; t < n_tensors; t += FIN_THREADS) {
    const float n = tnorms[t];
    acc += (double)n * (double)n;
    if (tnorm_out) tnorm_out[t] = n;
  }
  acc = rk_warp_sum(acc);
  if (lane == 0) s_red[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < FIN_WARPS; ++w) tot += s_red[w];
    const float total = (float)sqrt(tot);
    *reinterpret_cast<float*>(scratch + h->off_total) = total;
    if (total_out) *total_out = total;
  }
}

// 8 resident CTAs per SM (32 registers): the level-1 loop keeps one batch of loads per CTA in
// flight, so bytes in flight scale with residency (6 CTAs: 6.2 TB/s, 5 CTAs: 5.1 TB/s at 1e9).
__global__ void __launch_bounds__(NORM_THREADS, 8)
rk_l2norm_kernel(const char* __restrict__ plan, float grad_scale, char* __restrict__ scratch) {
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  __shared__ double s_red[2][NORM_WARPS];
  // level 2 (and behind it the step kernel) are launched as programmatic dependents: they set
  // themselves up while this kernel drains and read its output behind griddepcontrol.wait
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int n_chunks = h->n_chunks_norm;
  const int chunk = h->chunk_norm;
  const rk_tensor_desc* descs = reinterpret_cast<const rk_tensor_desc*>(plan + h->off_descs);
  const int32_t* gprefix = reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm);
  double* partials = reinterpret_cast<double*>(scratch + h->off_partials);

  const int32_t* prefix = rk_stage_prefix(gprefix, n_tensors + 1, s_prefix);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool stream_loads = h->total_numel > ((int64_t)16 << 20);   // > 64 MB of gradients

  int parity = 0;
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x, parity ^= 1) {
    const int t = rk_find_tensor(prefix, n_tensors, c);
    const int64_t numel = descs[t].numel;
    const int flags = descs[t].flags;
    const int64_t base = (int64_t)(c - prefix[t]) * chunk;
    const int64_t rem = numel - base;
    const int count = rem < chunk ? (int)rem : chunk;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (!(flags & RK_TENSOR_SKIP)) {
      const float* __restrict__ g = descs[t].g + base;
      if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
        const float4* __restrict__ g4 = reinterpret_cast<const float4*>(g);
        const int nvec = count >> 2;
        for (int i = tid; i < nvec; i += NORM_THREADS * NORM_UNROLL) {
          float4 x[NORM_UNROLL];
#pragma unroll
          for (int k = 0; k < NORM_UNROLL; ++k) {
            const int idx = i + k * NORM_THREADS;
            // lists that fit in L2 are read again by the step kernel right away: keep them
            // cached; larger ones are streamed (evict-first) so they do not thrash L2
            x[k] = idx >= nvec ? make_float4(0.f, 0.f, 0.f, 0.f)
                               : (stream_loads ? __ldcs(g4 + idx) : __ldg(g4 + idx));
          }
#pragma unroll
          for (int k = 0; k < NORM_UNROLL; ++k) {
            a0 = sq_acc(a0, x[k].x, grad_scale);
            a1 = sq_acc(a1, x[k].y, grad_scale);
            a2 = sq_acc(a2, x[k].z, grad_scale);
            a3 = sq_acc(a3, x[k].w, grad_scale);
          }
        }
        const int tail = nvec << 2;
        if (tail + tid < count) a0 = sq_acc(a0, g[tail + tid], grad_scale);
      } else {
        for (int i = tid; i < count; i += NORM_THREADS) a0 = sq_acc(a0, g[i], grad_scale);
      }
    }
    double s = (double)a0 + (double)a1 + ((double)a2 + (double)a3);
    s = rk_warp_sum(s);
    if (lane == 0) s_red[parity][warp] = s;
    __syncthreads();
    if (tid == 0) {
      double tot = 0.0;
#pragma unroll
      for (int w = 0; w < NORM_WARPS; ++w) tot += s_red[parity][w];
      partials[c] = tot;
    }
  }

}

// ---- host side ---------------------------------------------------------------------------
// Level 2 is launched with programmatic stream serialization: behind level 1 or the exchange
// kernel (both trigger early) it becomes resident while they drain; behind anything else the
// attribute has no effect and the launch is an ordinary stream-ordered one.
static int launch_finalize(const void* plan_dev, void* scratch_dev, float* total_norm_dev,
                           float* tensor_norms_dev, cudaStream_t s) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(FIN_THREADS);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  RK_CUDA(cudaLaunchKernelEx(&cfg, rk_l2norm_finalize_kernel, (const char*)plan_dev,
                             (char*)scratch_dev, total_norm_dev, tensor_norms_dev));
  g_rk_launches.fetch_add(1);
  return 0;
}

extern "C" int rk_l2norm_finalize(const void* plan_host, const void* plan_dev,
                                  void* scratch_dev, float* total_norm_dev,
                                  float* tensor_norms_dev, void* stream) {
  RK_REQUIRE(plan_host && plan_dev && scratch_dev, RK_ERR_INVALID,
             "rk_l2norm_finalize: null plan or scratch");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID, "rk_l2norm_finalize: bad plan magic");
  return launch_finalize(plan_dev, scratch_dev, total_norm_dev, tensor_norms_dev,
                         (cudaStream_t)stream);
}

extern "C" int rk_l2norm_multi(const void* plan_host, const void* plan_dev, float grad_scale,
                               void* scratch_dev, float* total_norm_dev,
                               float* tensor_norms_dev, void* stream) {
  RK_REQUIRE(plan_host && plan_dev && scratch_dev, RK_ERR_INVALID,
             "rk_l2norm_multi: null plan or scratch");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID, "rk_l2norm_multi: bad plan magic");
  int grid = 0;
  int rc = rk_persistent_grid((const void*)rk_l2norm_kernel, NORM_THREADS, 0,
                              h->n_chunks_norm, &grid);
  if (rc) return rc;
  rk_l2norm_kernel<<<grid, NORM_THREADS, 0, (cudaStream_t)stream>>>(
      (const char*)plan_dev, grad_scale, (char*)scratch_dev);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return launch_finalize(plan_dev, scratch_dev, total_norm_dev, tensor_norms_dev,
                         (cudaStream_t)stream);
}

// ---- the same norm for callers that have no scratch buffer to give ---------------------
// Scratch for rk_l2norm_multi_static lives in this module's own device image: one array per
// GPU that loads the library, held until the process exits.  In host code `g_static_scratch`
// is only a stand-in for the symbol, not a device address; the address is looked up with
// cudaGetSymbolAddress (with the target device current) and handed to the kernel.
static __device__ __align__(256) unsigned char g_static_scratch[RK_STATIC_SCRATCH_BYTES];

#define RK_MAX_DEVICES 64
static std::mutex g_static_mu;
static int g_static_resident[RK_MAX_DEVICES];  // persistent grid size, per device (set-up, once)

// The array's device address is looked up on EVERY call (a host-side table lookup in the CUDA
// runtime: no allocation, no stream work).  Caching it per device ordinal would hand a dangling
// pointer to the kernel after cudaDeviceReset or any other tear-down and re-creation of the
// device's primary context; the fresh lookup follows the context that is current now.
static int static_setup_current_device(void** scratch, int* resident) {
  int dev = -1;
  RK_CUDA(cudaGetDevice(&dev));
  RK_REQUIRE(dev >= 0 && dev < RK_MAX_DEVICES, RK_ERR_UNSUPPORTED,
             "rk_l2norm_multi_static: device ordinal %d not supported (max %d)", dev,
             RK_MAX_DEVICES - 1);
  void* p = nullptr;
  RK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
  RK_REQUIRE(p != nullptr, RK_ERR_STATE, "cudaGetSymbolAddress returned a null address");
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (g_static_resident[dev] == 0) {
    int sms = 0, per_sm = 0;
    RK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    RK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rk_l2norm_kernel,
                                                          NORM_THREADS, 0));
    RK_REQUIRE(sms > 0 && per_sm > 0, RK_ERR_STATE, "norm kernel does not fit on an SM");
    g_static_resident[dev] = sms * per_sm;
  }
  *scratch = p;
  *resident = g_static_resident[dev];
  return 0;
}

extern "C" int rk_l2norm_multi_static(const void* plan_host, const void* plan_dev,
                                      float grad_scale, float* total_norm_dev,
                                      float* tensor_norms_dev, void* stream) {
  RK_REQUIRE(plan_host && plan_dev, RK_ERR_INVALID, "rk_l2norm_multi_static: null plan");
  RK_REQUIRE(total_norm_dev, RK_ERR_INVALID,
             "rk_l2norm_multi_static: total_norm_dev is null (the result has nowhere to go)");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID,
             "rk_l2norm_multi_static: bad plan magic");
  RK_REQUIRE(h->scratch_bytes <= RK_STATIC_SCRATCH_BYTES, RK_ERR_TOO_LARGE,
             "rk_l2norm_multi_static: this tensor list needs %u bytes of scratch (%d norm "
             "chunks, %d tensors) but the built-in scratch holds %u; use rk_l2norm_multi "
             "with a caller-provided buffer",
             h->scratch_bytes, h->n_chunks_norm, h->n_tensors, (unsigned)RK_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  int resident = 0;
  int rc = static_setup_current_device(&scratch, &resident);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // The array still holds whatever the previous call (possibly with another plan, hence
  // another layout) left there.  Nothing in it has to start at zero any more: partials, norms
  // and the total are fully overwritten before they are read.  (The 128 bytes of the former
  // completion ticket are still cleared, so that the layout's one "must be zero" word is.)
  RK_CUDA(cudaMemsetAsync((char*)scratch + h->off_ticket, 0, 128, s));
  int grid = h->n_chunks_norm < resident ? (h->n_chunks_norm > 0 ? h->n_chunks_norm : 1)
                                         : resident;
  rk_l2norm_kernel<<<grid, NORM_THREADS, 0, s>>>((const char*)plan_dev, grad_scale,
                                                 (char*)scratch);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return launch_finalize(plan_dev, scratch, total_norm_dev, tensor_norms_dev, s);
}
