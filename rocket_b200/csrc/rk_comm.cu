// K6: gradient exchange fused with the 1/world_size scale and the norm's first level,
// over NVLink peer memory.  One kernel per bucket, per rank:
//
//   start barrier   every rank tells every peer "my gradients of this bucket are final"
//   reduce + write  rank r owns the norm chunks c with c % W == r.  For each it loads the
//                   chunk from all W ranks' gradient buffers (P2P loads over NVLink), sums
//                   them in rank order, scales by 1/W, accumulates the sum of squares, and
//                   stores the averaged chunk into all W buffers (P2P stores).  The chunk's
//                   sum of squares goes into every rank's norm scratch.
//   end barrier     "all my stores have landed" to every peer; wait for the same from them.
//
// Compared with copy-in / ncclAllReduce / copy-out + a separate norm pass: the gradients
// cross HBM once on the owner (read) and once everywhere (write), the wire carries the same
// 2(W-1)/W bytes per element as a ring, the 4 B/param norm pass disappears, and because each
// element is reduced by exactly one rank, in a fixed order, all replicas end up with
// bit-identical gradients and bit-identical norm partials.
//
// Memory: every rank owns one "arena" allocation [signals | norm scratch | gradients] that
// all peers have mapped (CUDA IPC, done by the host side); peers.arena[q] is rank q's arena
// as seen from this rank.  Cross-GPU flags carry a monotonically increasing epoch, so they
// never need resetting.
#include <stdlib.h>

#include "rk_common.cuh"
#include "rk_step_ops.cuh"

#define COMM_THREADS 256
#define COMM_WARPS (COMM_THREADS / 32)
#define COMM_UNROLL 2
#define COMM_UNROLL_MC 4

struct CommPeers {
  char* arena[RK_COMM_MAX_WORLD];
};

struct CommSignals {
  unsigned long long ready[RK_COMM_MAX_BUCKETS][RK_COMM_MAX_WORLD];
  unsigned long long done[RK_COMM_MAX_BUCKETS][RK_COMM_MAX_WORLD];
  unsigned int arrived[RK_COMM_MAX_BUCKETS];
  // how many times each bucket has been exchanged; kept on the device (not passed as a launch
  // argument) so that a kernel replayed from a CUDA graph still gets a fresh epoch every time
  unsigned long long epoch[RK_COMM_MAX_BUCKETS];
};
static_assert(sizeof(CommSignals) <= RK_COMM_SIGNAL_BYTES, "signal pad too small");

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {
  float4 r;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_peer4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVSwitch multicast ("multimem"): one load that returns the SUM over every rank's copy of the
// address, reduced inside the switch, and one store that the switch replicates to every rank.
__device__ __forceinline__ float4 mc_ld_reduce4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ float mc_ld_reduce1(const float* mc) {
  float r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];"
               : "=f"(r) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void mc_st4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void mc_st1(float* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(mc), "f"(v) : "memory");
}

// MC = false: P2P loads/stores on the peers' mapped arenas.
// MC = true : multimem.ld_reduce / multimem.st on the arenas' multicast mapping -- the switch
//             does the W-way sum and the W-way replication, so each GPU sends and receives
//             about G bytes instead of 2(W-1)/W * G (worth it from W = 4 up).
template <int W, bool MC, int UMC>
__global__ void __launch_bounds__(COMM_THREADS)
rk_allreduce_norm_kernel(const CommPeers peers, char* mc_arena, int rank,
                         const char* __restrict__ plan, unsigned long long scratch_off,
                         int chunk_lo, int chunk_hi, int bucket, float scale, int broadcast) {
  // broadcast != 0: all-reduce (every rank receives every averaged chunk).
  // broadcast == 0: reduce-scatter (only the owner keeps its averaged chunks; used by the
  //                 sharded optimizer step, which broadcasts parameters instead).
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  __shared__ double s_red[2][COMM_WARPS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  CommSignals* mine = reinterpret_cast<CommSignals*>(peers.arena[rank]);
  // every CTA reads the same value: it is only advanced by the last CTA, after all arrived
  const unsigned long long epoch = mine->epoch[bucket] + 1;

  // ---- start barrier ---------------------------------------------------------------
  if (blockIdx.x == 0 && tid < W)
    st_release_sys(&reinterpret_cast<CommSignals*>(peers.arena[tid])->ready[bucket][rank], epoch);

  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int chunk = h->chunk_norm;
  const rk_tensor_desc* descs = reinterpret_cast<const rk_tensor_desc*>(plan + h->off_descs);
  const int32_t* prefix = rk_stage_prefix(
      reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm), n_tensors + 1, s_prefix);

  if (tid < W) {
    while (ld_acquire_sys(&mine->ready[bucket][tid]) < epoch) { }
  }
  __syncthreads();

  // ---- owned chunks: c in [chunk_lo, chunk_hi), c % W == rank -------------------------
  int first = chunk_lo + ((rank - chunk_lo % W) + W) % W;
  int parity = 0;
  for (int c = first + (int)blockIdx.x * W; c < chunk_hi; c += (int)gridDim.x * W, parity ^= 1) {
    const int t = rk_find_tensor(prefix, n_tensors, c);
    const rk_tensor_desc d = descs[t];
    const int64_t base = (int64_t)(c - prefix[t]) * chunk;
    const int64_t rem = d.numel - base;
    const int count = rem < chunk ? (int)rem : chunk;
    if (!(d.flags & RK_TENSOR_SKIP)) {
      const unsigned long long off =
          (unsigned long long)(reinterpret_cast<char*>(d.g + base) - peers.arena[rank]);
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      if constexpr (MC) {
        float* m = reinterpret_cast<float*>(mc_arena + off);
        if ((off & 15) == 0) {
          const int nvec = count >> 2;
          for (int i0 = tid; i0 < nvec; i0 += COMM_THREADS * UMC) {
            float4 s[UMC];
#pragma unroll
            for (int k = 0; k < UMC; ++k) {
              const int i = i0 + k * COMM_THREADS;
              if (i < nvec) s[k] = mc_ld_reduce4(m + 4 * i);
            }
#pragma unroll
            for (int k = 0; k < UMC; ++k) {
              const int i = i0 + k * COMM_THREADS;
              if (i < nvec) {
                s[k].x = __fmul_rn(s[k].x, scale); s[k].y = __fmul_rn(s[k].y, scale);
                s[k].z = __fmul_rn(s[k].z, scale); s[k].w = __fmul_rn(s[k].w, scale);
                a0 = fmaf(s[k].x, s[k].x, a0); a1 = fmaf(s[k].y, s[k].y, a1);
                a2 = fmaf(s[k].z, s[k].z, a2); a3 = fmaf(s[k].w, s[k].w, a3);
                if (broadcast) mc_st4(m + 4 * i, s[k]);
                else st_peer4(reinterpret_cast<float*>(peers.arena[rank] + off) + 4 * i, s[k]);
              }
            }
          }
          const int i = (nvec << 2) + tid;
          if (i < count) {
            const float v = __fmul_rn(mc_ld_reduce1(m + i), scale);
            a0 = fmaf(v, v, a0);
            if (broadcast) mc_st1(m + i, v);
            else reinterpret_cast<float*>(peers.arena[rank] + off)[i] = v;
          }
        } else {
          for (int i = tid; i < count; i += COMM_THREADS) {
            const float v = __fmul_rn(mc_ld_reduce1(m + i), scale);
            a0 = fmaf(v, v, a0);
            if (broadcast) mc_st1(m + i, v);
            else reinterpret_cast<float*>(peers.arena[rank] + off)[i] = v;
          }
        }
      } else {
        float* g[W];
  #pragma unroll
        for (int q = 0; q < W; ++q) g[q] = reinterpret_cast<float*>(peers.arena[q] + off);
        if ((off & 15) == 0) {
          const int nvec = count >> 2;
          for (int i0 = tid; i0 < nvec; i0 += COMM_THREADS * COMM_UNROLL) {
            float4 x[COMM_UNROLL][W];
  #pragma unroll
            for (int k = 0; k < COMM_UNROLL; ++k) {
              const int i = i0 + k * COMM_THREADS;
              if (i < nvec) {
  #pragma unroll
                for (int q = 0; q < W; ++q) x[k][q] = ld_peer4(g[q] + 4 * i);
              }
            }
  #pragma unroll
            for (int k = 0; k < COMM_UNROLL; ++k) {
              const int i = i0 + k * COMM_THREADS;
              if (i < nvec) {
                float4 s = x[k][0];
  #pragma unroll
                for (int q = 1; q < W; ++q) {   // fixed rank order: identical on every owner
                  s.x = __fadd_rn(s.x, x[k][q].x);
                  s.y = __fadd_rn(s.y, x[k][q].y);
                  s.z = __fadd_rn(s.z, x[k][q].z);
                  s.w = __fadd_rn(s.w, x[k][q].w);
                }
                s.x = __fmul_rn(s.x, scale); s.y = __fmul_rn(s.y, scale);
                s.z = __fmul_rn(s.z, scale); s.w = __fmul_rn(s.w, scale);
                a0 = fmaf(s.x, s.x, a0); a1 = fmaf(s.y, s.y, a1);
                a2 = fmaf(s.z, s.z, a2); a3 = fmaf(s.w, s.w, a3);
  #pragma unroll
                for (int q = 0; q < W; ++q)
                  if (broadcast || q == rank) st_peer4(g[q] + 4 * i, s);
              }
            }
          }
          const int i = (nvec << 2) + tid;
          if (i < count) {
            float s = g[0][i];
  #pragma unroll
            for (int q = 1; q < W; ++q) s = __fadd_rn(s, g[q][i]);
            s = __fmul_rn(s, scale);
            a0 = fmaf(s, s, a0);
  #pragma unroll
            for (int q = 0; q < W; ++q)
              if (broadcast || q == rank) g[q][i] = s;
          }
        } else {
          for (int i = tid; i < count; i += COMM_THREADS) {
            float s = g[0][i];
  #pragma unroll
            for (int q = 1; q < W; ++q) s = __fadd_rn(s, g[q][i]);
            s = __fmul_rn(s, scale);
            a0 = fmaf(s, s, a0);
  #pragma unroll
            for (int q = 0; q < W; ++q)
              if (broadcast || q == rank) g[q][i] = s;
          }
        }
      }
      double sd = (double)a0 + (double)a1 + ((double)a2 + (double)a3);
      sd = rk_warp_sum(sd);
      if (lane == 0) s_red[parity][warp] = sd;
    } else if (lane == 0) {
      s_red[parity][warp] = 0.0;
    }
    __syncthreads();
    if (tid < W) {   // one thread per destination rank publishes this chunk's partial
      double tot = 0.0;
#pragma unroll
      for (int w = 0; w < COMM_WARPS; ++w) tot += s_red[parity][w];
      double* partials =
          reinterpret_cast<double*>(peers.arena[tid] + scratch_off + h->off_partials);
      partials[c] = tot;
    }
  }

  // ---- end barrier ---------------------------------------------------------------------
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const unsigned int prev = atomicAdd(&mine->arrived[bucket], 1u);
    if (prev == gridDim.x - 1) {
      mine->arrived[bucket] = 0u;
      mine->epoch[bucket] = epoch;
      __threadfence_system();
      for (int q = 0; q < W; ++q)
        st_release_sys(&reinterpret_cast<CommSignals*>(peers.arena[q])->done[bucket][rank], epoch);
      for (int q = 0; q < W; ++q)
        while (ld_acquire_sys(&mine->done[bucket][q]) < epoch) { }
      __threadfence_system();
    }
  }
}

struct rk_comm {
  int rank, world, device, sms;
  CommPeers peers;
  char* mc_arena;   // multicast mapping of all arenas, or nullptr
  unsigned long long scratch_off;
};

extern "C" int rk_comm_create(rk_comm** out, int rank, int world, void* const* arenas,
                              void* multicast_arena, uint64_t scratch_offset) {
  RK_REQUIRE(out && arenas, RK_ERR_INVALID, "rk_comm_create: null argument");
  RK_REQUIRE(world == 2 || world == 4 || world == 8, RK_ERR_UNSUPPORTED,
             "rk_comm_create: world size %d (supported: 2, 4, 8)", world);
  RK_REQUIRE(rank >= 0 && rank < world, RK_ERR_INVALID, "rk_comm_create: rank %d of %d", rank, world);
  RK_REQUIRE(scratch_offset >= RK_COMM_SIGNAL_BYTES && scratch_offset % 256 == 0, RK_ERR_INVALID,
             "rk_comm_create: scratch must follow the %d-byte signal pad, 256-byte aligned",
             RK_COMM_SIGNAL_BYTES);
  rk_comm* c = new rk_comm();
  c->rank = rank;
  c->world = world;
  c->scratch_off = scratch_offset;
  c->mc_arena = (char*)multicast_arena;
  for (int q = 0; q < RK_COMM_MAX_WORLD; ++q) c->peers.arena[q] = nullptr;
  for (int q = 0; q < world; ++q) {
    if (!arenas[q]) {
      delete c;
      RK_REQUIRE(false, RK_ERR_INVALID, "rk_comm_create: arena of rank %d is null", q);
    }
    c->peers.arena[q] = (char*)arenas[q];
  }
  cudaError_t e = cudaGetDevice(&c->device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, c->device);
  if (e != cudaSuccess) {
    rk_set_error("rk_comm_create: %s", cudaGetErrorString(e));
    delete c;
    return (int)e;
  }
  *out = c;
  return 0;
}

static int launch_exchange(rk_comm* c, const void* plan_host, const void* plan_dev, int chunk_lo,
                           int chunk_hi, int bucket, int max_ctas, int use_multicast,
                           int broadcast, void* stream) {
  RK_REQUIRE(c && plan_host && plan_dev, RK_ERR_INVALID, "rk_comm_allreduce_norm: null argument");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID, "rk_comm_allreduce_norm: bad plan magic");
  RK_REQUIRE(bucket >= 0 && bucket < RK_COMM_MAX_BUCKETS, RK_ERR_TOO_LARGE,
             "rk_comm_allreduce_norm: bucket %d outside [0,%d)", bucket, RK_COMM_MAX_BUCKETS);
  RK_REQUIRE(0 <= chunk_lo && chunk_lo <= chunk_hi && chunk_hi <= h->n_chunks_norm, RK_ERR_INVALID,
             "rk_comm_allreduce_norm: chunk range [%d,%d) outside [0,%d]", chunk_lo, chunk_hi,
             h->n_chunks_norm);
  const int W = c->world;
  const int owned = (chunk_hi - chunk_lo + W - 1) / W;
  // reduce-scatter keeps its stores local and is limited by loads in flight: two CTAs per SM
  // (1e9 params, 2 GPUs: 4.09 -> 3.27 ms; all-reduce does not gain)
  int grid = max_ctas > 0 ? max_ctas : (broadcast ? c->sms : 2 * c->sms);
  if (grid > owned) grid = owned;
  if (grid < 1) grid = 1;   // still takes part in both barriers
  const float scale = 1.0f / (float)W;
  cudaStream_t s = (cudaStream_t)stream;
  const bool mc = c->mc_arena != nullptr && use_multicast != 0;
  // multimem loads a thread keeps in flight (tuning: RK_COMM_MC_UNROLL = 2 | 4 | 8)
  int umc = COMM_UNROLL_MC;
  if (mc) {
    const char* e = getenv("RK_COMM_MC_UNROLL");
    if (e) umc = atoi(e);
  }
#define RK_LAUNCH(WW, MC, U)                                                                 \
  rk_allreduce_norm_kernel<WW, MC, U><<<grid, COMM_THREADS, 0, s>>>(c->peers, c->mc_arena,   \
      c->rank, (const char*)plan_dev, c->scratch_off, chunk_lo, chunk_hi, bucket, scale,    \
      broadcast)
#define RK_LAUNCH_W(WW)                                                                      \
  if (!mc) RK_LAUNCH(WW, false, 4);                                                          \
  else if (umc == 2) RK_LAUNCH(WW, true, 2);                                                 \
  else if (umc == 8) RK_LAUNCH(WW, true, 8);                                                 \
  else RK_LAUNCH(WW, true, 4)
  switch (W) {
    case 2: RK_LAUNCH_W(2); break;
    case 4: RK_LAUNCH_W(4); break;
    case 8: RK_LAUNCH_W(8); break;
  }
#undef RK_LAUNCH_W
#undef RK_LAUNCH
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
}

extern "C" int rk_comm_allreduce_norm(rk_comm* c, const void* plan_host, const void* plan_dev,
                                      int chunk_lo, int chunk_hi, int bucket, int max_ctas,
                                      int use_multicast, void* stream) {
  return launch_exchange(c, plan_host, plan_dev, chunk_lo, chunk_hi, bucket, max_ctas,
                         use_multicast, /*broadcast=*/1, stream);
}

extern "C" int rk_comm_reduce_scatter_norm(rk_comm* c, const void* plan_host,
                                           const void* plan_dev, int chunk_lo, int chunk_hi,
                                           int bucket, int max_ctas, int use_multicast,
                                           void* stream) {
  return launch_exchange(c, plan_host, plan_dev, chunk_lo, chunk_hi, bucket, max_ctas,
                         use_multicast, /*broadcast=*/0, stream);
}

// ---- sharded optimizer step ------------------------------------------------------------
// After rk_comm_reduce_scatter_norm every rank holds the averaged gradient of the norm chunks
// it owns (c % W == rank).  This kernel walks ALL chunks: an owned chunk gets the fused
// AdamW / SGD update on the local p, g, m, v and its new parameters are written into every
// rank's parameter buffer (P2P stores, or one multimem.st that the switch replicates); any
// other chunk only has its local gradient zero-filled.  Optimizer work and moment traffic per
// GPU shrink by W; per element the arithmetic is the replicated step's, so the result is
// bit-identical to data parallelism.  One end barrier: nobody's next forward may start
// before all owners' parameters have landed.
template <class Op, class Table, int W, bool MC>
__global__ void __launch_bounds__(COMM_THREADS, 3)
rk_sharded_step_kernel(const CommPeers peers, char* mc_arena, int rank,
                       const char* __restrict__ plan, const Table* __restrict__ dtable,
                       float* __restrict__ shard_m, float* __restrict__ shard_v,
                       const float* __restrict__ total_norm, float max_norm, float grad_scale,
                       int bucket) {
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  const int tid = threadIdx.x;
  CommSignals* mine = reinterpret_cast<CommSignals*>(peers.arena[rank]);
  const unsigned long long epoch = mine->epoch[bucket] + 1;
  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int n_chunks = h->n_chunks_norm;
  const int chunk = h->chunk_norm;
  const rk_tensor_desc* descs = reinterpret_cast<const rk_tensor_desc*>(plan + h->off_descs);
  const int32_t* prefix = rk_stage_prefix(
      reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm), n_tensors + 1, s_prefix);
  Prolog pr;
  pr.scale = grad_scale;                 // 1/loss_scale under fp16 (the exchange already averaged)
  pr.has_scale = grad_scale != 1.0f;
  pr.has_clip = total_norm != nullptr;
  pr.coef = pr.has_clip ? rk_clip_coef(__ldg(total_norm), max_norm) : 1.0f;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

  // A CTA takes whole rows of W consecutive chunks: one it owns (heavy: update + W-way store)
  // and W-1 it only zero-fills.  Striding single chunks by gridDim would hand all the owned
  // chunks to 1/W of the CTAs whenever gridDim is a multiple of W.
  const int rows = (n_chunks + W - 1) / W;
  for (int k = blockIdx.x; k < rows; k += gridDim.x)
  for (int q = 0; q < W; ++q) {
    const int c = k * W + q;
    if (c >= n_chunks) break;
    const int t = rk_find_tensor(prefix, n_tensors, c);
    const rk_tensor_desc d = descs[t];
    Op op;
    op.h = dtable->g[d.group];
    if ((d.flags & RK_TENSOR_SKIP) || (op.h.flags & GROUP_SKIP)) continue;
    const int64_t base = (int64_t)(c - prefix[t]) * chunk;
    const int64_t rem = d.numel - base;
    const int count = rem < chunk ? (int)rem : chunk;
    float* P = d.p + base;
    float* G = d.g + base;
    const bool vec = (d.flags & RK_TENSOR_VEC4) != 0;
    if (q != rank) {                         // not mine: only clear my local gradient
      if (vec) {
        const int nvec = count >> 2;
        for (int i = tid; i < nvec; i += COMM_THREADS) stg4<0>(G + 4 * i, zero4);
        const int i = (nvec << 2) + tid;
        if (i < count) stg1(G + i, 0.f);
      } else {
        for (int i = tid; i < count; i += COMM_THREADS) stg1(G + i, 0.f);
      }
      continue;
    }
    // moments: compact per-rank buffers (owned chunk c lives in slot c / W = this row), or the
    // plan's full-size per-tensor buffers
    const int64_t slot = (int64_t)k * chunk;
    float* M = shard_m ? shard_m + slot : d.m + base;
    float* V = !Op::kHasV ? nullptr : (shard_v ? shard_v + slot : d.v + base);
    const bool has_m = shard_m != nullptr || d.m != nullptr;
    const unsigned long long poff =
        (unsigned long long)(reinterpret_cast<char*>(P) - peers.arena[rank]);
    auto put_p4 = [&](int i, const float4& v) {
      if constexpr (MC) {
        mc_st4(reinterpret_cast<float*>(mc_arena + poff) + 4 * i, v);
      } else {
#pragma unroll
        for (int q = 0; q < W; ++q)
          st_peer4(reinterpret_cast<float*>(peers.arena[q] + poff) + 4 * i, v);
      }
    };
    auto put_p1 = [&](int i, float v) {
      if constexpr (MC) {
        mc_st1(reinterpret_cast<float*>(mc_arena + poff) + i, v);
      } else {
#pragma unroll
        for (int q = 0; q < W; ++q) reinterpret_cast<float*>(peers.arena[q] + poff)[i] = v;
      }
    };
    auto one = [&](int i) {
      float p = ldg1(P + i), m = has_m ? ldg1(M + i) : 0.f, v = Op::kHasV ? ldg1(V + i) : 0.f;
      op(p, prolog_g(ldg1(G + i), pr), m, v);
      put_p1(i, p);
      if (has_m) stg1(M + i, m);
      if (Op::kHasV) stg1(V + i, v);
      stg1(G + i, 0.f);
    };
    if (vec) {
      const int nvec = count >> 2;
      for (int i = tid; i < nvec; i += COMM_THREADS) {
        float4 p4 = ldg4<1>(P + 4 * i), g4 = ldg4<1>(G + 4 * i), m4 = zero4, v4 = zero4;
        if (has_m) m4 = ldg4<1>(M + 4 * i);
        if (Op::kHasV) v4 = ldg4<1>(V + 4 * i);
        op(p4.x, prolog_g(g4.x, pr), m4.x, v4.x);
        op(p4.y, prolog_g(g4.y, pr), m4.y, v4.y);
        op(p4.z, prolog_g(g4.z, pr), m4.z, v4.z);
        op(p4.w, prolog_g(g4.w, pr), m4.w, v4.w);
        put_p4(i, p4);
        if (has_m) stg4<0>(M + 4 * i, m4);
        if (Op::kHasV) stg4<0>(V + 4 * i, v4);
        stg4<0>(G + 4 * i, zero4);
      }
      const int i = (nvec << 2) + tid;
      if (i < count) one(i);
    } else {
      for (int i = tid; i < count; i += COMM_THREADS) one(i);
    }
  }

  // ---- end barrier: all owners' parameters are in place everywhere ------------------------
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const unsigned int prev = atomicAdd(&mine->arrived[bucket], 1u);
    if (prev == gridDim.x - 1) {
      mine->arrived[bucket] = 0u;
      mine->epoch[bucket] = epoch;
      __threadfence_system();
      for (int q = 0; q < W; ++q)
        st_release_sys(&reinterpret_cast<CommSignals*>(peers.arena[q])->done[bucket][rank], epoch);
      for (int q = 0; q < W; ++q)
        while (ld_acquire_sys(&mine->done[bucket][q]) < epoch) { }
      __threadfence_system();
    }
  }
}

template <class Op, class Table>
static int launch_sharded(rk_comm* c, const void* plan_host, const void* plan_dev,
                          const void* table_dev, float* shard_m, float* shard_v,
                          const float* total_norm_dev, float max_norm, float grad_scale,
                          int max_ctas, int use_multicast, void* stream) {
  RK_REQUIRE(c && plan_host && plan_dev && table_dev, RK_ERR_INVALID,
             "rk_comm_sharded_step: null argument");
  RK_REQUIRE(!Op::kHasV || (shard_m == nullptr) == (shard_v == nullptr), RK_ERR_INVALID,
             "rk_comm_sharded_step: shard_m and shard_v must both be given or both be NULL");
  RK_REQUIRE((((uintptr_t)shard_m | (uintptr_t)shard_v) & 15) == 0, RK_ERR_INVALID,
             "rk_comm_sharded_step: moment shards must be 16-byte aligned");
  const rk_plan_header* h = (const rk_plan_header*)plan_host;
  RK_REQUIRE(h->magic == RK_PLAN_MAGIC, RK_ERR_INVALID, "rk_comm_sharded_step: bad plan magic");
  int grid = max_ctas > 0 ? max_ctas : 3 * c->sms;
  const int rows = (h->n_chunks_norm + c->world - 1) / c->world;
  if (grid > rows) grid = rows;
  if (grid < 1) grid = 1;
  const int bucket = RK_COMM_MAX_BUCKETS - 1;   // this kernel's own barrier slot
  const bool mc = c->mc_arena != nullptr && use_multicast != 0;
  cudaStream_t s = (cudaStream_t)stream;
#define RK_LAUNCH(WW, MC)                                                                     \
  rk_sharded_step_kernel<Op, Table, WW, MC><<<grid, COMM_THREADS, 0, s>>>(                   \
      c->peers, c->mc_arena, c->rank, (const char*)plan_dev, (const Table*)table_dev,         \
      shard_m, shard_v, total_norm_dev, max_norm, grad_scale, bucket)
  switch (c->world) {
    case 2: if (mc) RK_LAUNCH(2, true); else RK_LAUNCH(2, false); break;
    case 4: if (mc) RK_LAUNCH(4, true); else RK_LAUNCH(4, false); break;
    case 8: if (mc) RK_LAUNCH(8, true); else RK_LAUNCH(8, false); break;
  }
#undef RK_LAUNCH
  RK_CUDA(cudaGetLastError());
  g_rk_launches.fetch_add(1);
  return 0;
}

extern "C" int rk_comm_sharded_step(rk_comm* c, const void* plan_host, const void* plan_dev,
                                    int optimizer, const void* table_dev, float* shard_m,
                                    float* shard_v, const float* total_norm_dev, float max_norm,
                                    float grad_scale, int max_ctas, int use_multicast,
                                    void* stream) {
  if (optimizer == 0)
    return launch_sharded<AdamOp, AdamTable>(c, plan_host, plan_dev, table_dev, shard_m, shard_v,
                                             total_norm_dev, max_norm, grad_scale, max_ctas, use_multicast,
                                             stream);
  if (optimizer == 1)
    return launch_sharded<SgdOp, SgdTable>(c, plan_host, plan_dev, table_dev, shard_m, shard_v,
                                           total_norm_dev, max_norm, grad_scale, max_ctas, use_multicast,
                                           stream);
  RK_REQUIRE(false, RK_ERR_INVALID, "rk_comm_sharded_step: optimizer %d (0 = Adam(W), 1 = SGD)",
             optimizer);
}

extern "C" int rk_comm_destroy(rk_comm* c) {
  delete c;
  return 0;
}
