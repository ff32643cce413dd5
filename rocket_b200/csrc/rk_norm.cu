// K1: deterministic multi-tensor gradient L2 norm (HBM-bound: 4 B read per element).
//
// Level 1: every norm chunk (plan.chunk_norm elements of one tensor) -> one double
//          partial, written to scratch in chunk order.
// Level 2: the CTA that finishes last (ticket counter) reduces, in a fixed order,
//          partials -> per-tensor norms -> total norm.  No float atomics anywhere, so
//          the result does not depend on CTA scheduling and is bit-identical on every
//          rank that holds the same gradients.
#include <mutex>

#include "rk_common.cuh"

#define NORM_THREADS 256
#define NORM_WARPS (NORM_THREADS / 32)
#define NORM_UNROLL 4
#define NORM_BIG_TENSOR 1024  // tensors with more chunks than this are reduced block-wide
#define NORM_THREAD_TENSOR 8  // tensors with at most this many chunks are summed by one thread

__device__ __forceinline__ float sq_acc(float acc, float x, float s) {
  float y = x * s;
  return fmaf(y, y, acc);
}

// ---- level 2: partials -> per-tensor norms -> total norm, by ONE CTA, in a fixed order ----
// `gprefix`: the norm-chunk prefix table, in shared memory whenever the caller could stage it
// (a global-memory table costs one L2 round trip per tensor in loops that branch on it: that
// was 15 of the 27 us this kernel took on a ResNet-18, profiles/r2_norm_notes.md).
__device__ __forceinline__ void rk_norm_level2(const rk_plan_header* h,
                                               const int32_t* gprefix,
                                               const double* partials, float* tnorms,
                                               char* scratch, float* total_out,
                                               float* tnorm_out, unsigned int* ticket,
                                               double (*s_red)[NORM_WARPS]) {
  const int n_tensors = h->n_tensors;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // ---- level 2a: per-tensor norms ---------------------------------------------------
  // Tensors of at most NORM_THREAD_TENSOR chunks: one THREAD each, summing its partials in
  // chunk order, four loads in flight.  A list is mostly such tensors (biases, norms, small
  // matrices), and one warp per tensor made this loop a chain of per-tensor round trips: 186 us
  // for the 1185 tensors of the configs[4] list (tools/diag_norm_split.py).
  int any_big = 0, any_mid = 0;
  for (int t = tid; t < n_tensors; t += NORM_THREADS) {
    const int c0 = gprefix[t], c1 = gprefix[t + 1];
    if (c1 - c0 > NORM_THREAD_TENSOR) { any_mid = 1; continue; }
    double s = 0.0;
    int c = c0;
    for (; c + 3 < c1; c += 4) {
      const double x0 = __ldcg(partials + c), x1 = __ldcg(partials + c + 1);
      const double x2 = __ldcg(partials + c + 2), x3 = __ldcg(partials + c + 3);
      s += x0; s += x1; s += x2; s += x3;
    }
    for (; c < c1; ++c) s += __ldcg(partials + c);
    tnorms[t] = (float)sqrt(s);
  }
  // ... up to NORM_BIG_TENSOR chunks: one warp each
  any_mid = __syncthreads_or(any_mid);
  for (int t = warp; any_mid && t < n_tensors; t += NORM_WARPS) {
    const int c0 = gprefix[t], c1 = gprefix[t + 1];
    if (c1 - c0 <= NORM_THREAD_TENSOR) continue;
    if (c1 - c0 > NORM_BIG_TENSOR) { any_big = 1; continue; }
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int c = c0 + lane;
    for (; c + 96 < c1; c += 128) {
      s0 += __ldcg(partials + c);
      s1 += __ldcg(partials + c + 32);
      s2 += __ldcg(partials + c + 64);
      s3 += __ldcg(partials + c + 96);
    }
    for (; c < c1; c += 32) s0 += __ldcg(partials + c);
    const double s = rk_warp_sum((s0 + s1) + (s2 + s3));
    if (lane == 0) tnorms[t] = (float)sqrt(s);
  }
  // ---- level 2b: big tensors, whole CTA each ---------------------------------------
  int par = 0;
  any_big = __syncthreads_or(any_big);
  for (int t = 0; any_big && t < n_tensors; ++t) {
    const int c0 = gprefix[t], c1 = gprefix[t + 1];
    if (c1 - c0 <= NORM_BIG_TENSOR) continue;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int c = c0 + tid;
    for (; c + 3 * NORM_THREADS < c1; c += 4 * NORM_THREADS) {
      s0 += __ldcg(partials + c);
      s1 += __ldcg(partials + c + NORM_THREADS);
      s2 += __ldcg(partials + c + 2 * NORM_THREADS);
      s3 += __ldcg(partials + c + 3 * NORM_THREADS);
    }
    for (; c < c1; c += NORM_THREADS) s0 += __ldcg(partials + c);
    double s = rk_warp_sum((s0 + s1) + (s2 + s3));
    if (lane == 0) s_red[par][warp] = s;
    __syncthreads();
    if (tid == 0) {
      double tot = 0.0;
#pragma unroll
      for (int w = 0; w < NORM_WARPS; ++w) tot += s_red[par][w];
      tnorms[t] = (float)sqrt(tot);
    }
    par ^= 1;
  }
  __syncthreads();  // tnorms written by this CTA are visible to this CTA
  // ---- level 2c: norm of the per-tensor norms (clip_grad.py:106-107) -----------------
  double acc = 0.0;
  for (int t = tid; t < n_tensors; t += NORM_THREADS) {
    const float n = tnorms[t];
    acc += (double)n * (double)n;
    if (tnorm_out) tnorm_out[t] = n;
  }
  acc = rk_warp_sum(acc);
  if (lane == 0) s_red[par][warp] = acc;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < NORM_WARPS; ++w) tot += s_red[par][w];
    const float total = (float)sqrt(tot);
    *reinterpret_cast<float*>(scratch + h->off_total) = total;
    if (total_out) *total_out = total;
    *ticket = 0u;  // ready for the next launch
  }
}

// 8 resident CTAs per SM (32 registers): the level-1 loop keeps one batch of loads per CTA in
// flight, so bytes in flight scale with residency (6 CTAs: 6.2 TB/s, 5 CTAs: 5.1 TB/s at 1e9).
__global__ void __launch_bounds__(NORM_THREADS, 8)
rk_l2norm_kernel(const char* __restrict__ plan, float grad_scale, char* __restrict__ scratch,
                 float* __restrict__ total_out, float* __restrict__ tnorm_out) {
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  __shared__ double s_red[2][NORM_WARPS];
  __shared__ int s_last;
  // a step kernel launched as a programmatic dependent may set itself up (and prefetch) while
  // this kernel runs; it reads the norm only behind its griddepcontrol.wait
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int n_chunks = h->n_chunks_norm;
  const int chunk = h->chunk_norm;
  const rk_tensor_desc* descs = reinterpret_cast<const rk_tensor_desc*>(plan + h->off_descs);
  const int32_t* gprefix = reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm);
  double* partials = reinterpret_cast<double*>(scratch + h->off_partials);
  float* tnorms = reinterpret_cast<float*>(scratch + h->off_tnorms);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + h->off_ticket);

  const int32_t* prefix = rk_stage_prefix(gprefix, n_tensors + 1, s_prefix);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool stream_loads = h->total_numel > ((int64_t)16 << 20);   // > 64 MB of gradients

  // Which tensor a chunk belongs to, where it starts and how long it is.  The lookup for the
  // CTA's next chunk (binary search in shared memory + one descriptor read from L2) is issued
  // while the current chunk's loads are in flight, so it costs no time between chunks.
  struct Chunk { const float* g; int count; };
  auto lookup = [&](int c) {
    Chunk w;
    w.g = nullptr;
    w.count = 0;
    if (c < n_chunks) {
      const int t = rk_find_tensor(prefix, n_tensors, c);
      const int64_t base = (int64_t)(c - prefix[t]) * chunk;
      const int64_t rem = descs[t].numel - base;
      w.count = rem < chunk ? (int)rem : chunk;
      w.g = (descs[t].flags & RK_TENSOR_SKIP) ? nullptr : descs[t].g + base;
    }
    return w;
  };
  auto load_batch = [&](const float4* __restrict__ g4, int i, int nvec, float4 (&x)[NORM_UNROLL]) {
#pragma unroll
    for (int k = 0; k < NORM_UNROLL; ++k) {
      const int idx = i + k * NORM_THREADS;
      // lists that fit in L2 are read again by the step kernel right away: keep them
      // cached; larger ones are streamed (evict-first) so they do not thrash L2
      x[k] = idx >= nvec ? make_float4(0.f, 0.f, 0.f, 0.f)
                         : (stream_loads ? __ldcs(g4 + idx) : __ldg(g4 + idx));
    }
  };

  int parity = 0;
  Chunk cur = lookup(blockIdx.x);
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x, parity ^= 1) {
    Chunk nxt;
    bool have_next = false;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (cur.g != nullptr) {
      const float* __restrict__ g = cur.g;
      const int count = cur.count;
      if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
        const float4* __restrict__ g4 = reinterpret_cast<const float4*>(g);
        const int nvec = count >> 2;
        for (int i = tid; i < nvec; i += NORM_THREADS * NORM_UNROLL) {
          float4 x[NORM_UNROLL];
          load_batch(g4, i, nvec, x);
          if (!have_next) {                 // behind the first batch of loads
            nxt = lookup(c + gridDim.x);
            have_next = true;
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
    if (!have_next) nxt = lookup(c + gridDim.x);
    cur = nxt;
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

  // ---- ticket: who is last? ------------------------------------------------------
  if (tid == 0) {
    __threadfence();
    unsigned int prev = atomicAdd(ticket, 1u);
    s_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  rk_norm_level2(h, prefix, partials, tnorms, scratch, total_out, tnorm_out, ticket, s_red);
}

// Level 2 on its own: for partials produced elsewhere (the fused all-reduce kernel writes
// them, rk_comm.cu).  One CTA.
__global__ void __launch_bounds__(NORM_THREADS)
rk_l2norm_finalize_kernel(const char* __restrict__ plan, char* __restrict__ scratch,
                          float* __restrict__ total_out, float* __restrict__ tnorm_out) {
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  __shared__ double s_red[2][NORM_WARPS];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int32_t* gprefix = rk_stage_prefix(
      reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm), h->n_tensors + 1, s_prefix);
  rk_norm_level2(h, gprefix, reinterpret_cast<const double*>(scratch + h->off_partials),
                 reinterpret_cast<float*>(scratch + h->off_tnorms), scratch, total_out,
                 tnorm_out, reinterpret_cast<unsigned int*>(scratch + h->off_ticket), s_red);
}

extern "C" int rk_l2norm_finalize(const void* plan_host, const void* plan_dev,
                                  void* scratch_dev, float* total_norm_dev,
                                  float* tensor_norms_dev, void* stream) {
  RK_REQUIRE(plan_host && plan_dev && scratch_dev, RK_ERR_INVALID,
             "rk_l2norm_finalize: null plan or scratch");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID, "rk_l2norm_finalize: bad plan magic");
  rk_l2norm_finalize_kernel<<<1, NORM_THREADS, 0, (cudaStream_t)stream>>>(
      (const char*)plan_dev, (char*)scratch_dev, total_norm_dev, tensor_norms_dev);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
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
      (const char*)plan_dev, grad_scale, (char*)scratch_dev, total_norm_dev, tensor_norms_dev);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
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
  // another layout) left there.  Only the ticket has to start at zero; partials and norms
  // are fully overwritten before they are read.
  RK_CUDA(cudaMemsetAsync((char*)scratch + h->off_ticket, 0, 128, s));
  int grid = h->n_chunks_norm < resident ? (h->n_chunks_norm > 0 ? h->n_chunks_norm : 1)
                                         : resident;
  rk_l2norm_kernel<<<grid, NORM_THREADS, 0, s>>>((const char*)plan_dev, grad_scale,
                                                 (char*)scratch, total_norm_dev,
                                                 tensor_norms_dev);
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
}
