// K2 / K3, bulk-copy form: the same fused optimizer step as rk_step_kernel, with the HBM
// traffic moved by the copy engine path (cp.async.bulk, the non-tensor form of TMA) instead of
// per-thread LDG/STG.
//
// One persistent CTA per SM.  Warp 0 is the producer: it walks the CTA's share of the step
// chunks, and for each tile issues up to four bulk loads (p, g, m, v: contiguous 16-byte
// aligned runs) into one stage of a shared-memory ring, completion counted in bytes on the
// stage's `full` mbarrier.  The other warps are consumers: they wait for `full`, update the
// tile in place in shared memory with exactly the element code of the LDG kernel
// (rk_step_ops.cuh, so results stay bit-identical), make their writes visible to the async
// proxy and arrive on the stage's `done` mbarrier.  The producer then issues the bulk stores
// (p, m, v and the zeroed g) from the same stage and recycles it once the stores have read
// shared memory (cp.async.bulk.wait_group.read).
//
// Why: bytes in flight no longer cost registers (a 6-stage ring keeps ~128 KB of loads per SM
// outstanding behind 16 compute warps with ~40 registers each), loads and stores reach the
// memory system as whole contiguous tiles (fewer DRAM read/write turnarounds than eight
// interleaved per-thread streams), and the LSU issues one instruction per tile and stream
// instead of one per 16 bytes.
//
// Tiles that cannot be moved in bulk (tensor not 16-byte aligned) are handed to the consumers
// as "direct" items: they process them straight from global memory like the LDG kernel.  A
// tile's last count % 4 elements are always done that way.
#pragma once
#include "rk_step_ops.cuh"

#define BULK_MAX_STAGES 12
#define BULK_SMEM_DATA (192 * 1024)   // ring bytes; + 16 KB prefix table + meta < 227 KB

struct BulkItem {            // what the producer tells the consumers about one stage
  float* P;
  float* G;
  float* M;
  float* V;
  int count;                 // elements in this tile
  int nbulk;                 // leading elements that went through shared memory (multiple of 4)
  int group;
  int mode;                  // 0 = tile, 1 = end of stream
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes,
                                          uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Programmatic dependent launch: when this kernel is launched with the programmatic-stream-
// serialization attribute right behind a kernel that triggers early (the norm kernels do), its
// CTAs may start while that kernel is still finishing.  Everything that depends on the
// previous kernel's output (the norm) or overwrites what it reads (the gradients) comes after
// this wait; the producer's first bulk loads do not, and overlap the previous kernel's tail.
__device__ __forceinline__ void grid_dependency_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// CW consumer warps + 1 producer warp.  `tile_log2`: elements per stage and stream (a power of
// two <= 4096, so it divides the plan's norm chunk), `stages` ring slots; chosen by the host.
template <class Op, class Table, int CW>
__global__ void __launch_bounds__(32 * (CW + 1), 1)
rk_step_bulk_kernel(const char* __restrict__ plan, const __grid_constant__ Table table,
                    const Table* __restrict__ dtable, const float* __restrict__ total_norm,
                    float max_norm, float grad_scale, int zero_grads, int tile_log2,
                    int stages) {
  extern __shared__ __align__(128) unsigned char s_ring[];
  __shared__ int32_t s_prefix[RK_SMEM_PREFIX_CAP];
  __shared__ __align__(8) uint64_t s_full[BULK_MAX_STAGES];
  __shared__ __align__(8) uint64_t s_done[BULK_MAX_STAGES];
  __shared__ BulkItem s_item[BULK_MAX_STAGES];

  // Work items are tiles of the NORM chunking (chunks of >= 4096 elements, cut into `tile`
  // element pieces): large enough that the per-tile hand-off between producer and consumers
  // is amortised, whatever the step chunking of the LDG kernels is.
  const rk_plan_header* h = reinterpret_cast<const rk_plan_header*>(plan);
  const int n_tensors = h->n_tensors;
  const int tile = 1 << tile_log2;
  const int chunk = h->chunk_norm;
  const int sub_log2 = __ffs(chunk) - 1 - tile_log2;     // log2(tiles per chunk)
  const int sub_mask = (1 << sub_log2) - 1;
  const int n_items = h->n_chunks_norm << sub_log2;        // host checked: fits in int32
  const rk_tensor_desc* descs = reinterpret_cast<const rk_tensor_desc*>(plan + h->off_descs);
  const int32_t* prefix = rk_stage_prefix(
      reinterpret_cast<const int32_t*>(plan + h->off_prefix_norm), n_tensors + 1, s_prefix);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t stage_bytes = (uint32_t)tile * 16u;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&s_done[s], CW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_smem();
  }
  __syncthreads();

  if (warp == 0) {
    // ================= producer (one lane) =================
    if (lane != 0) return;
    int it = 0;                 // items enqueued so far ...
    int it_s = 0;               // ... the ring slot the next one takes
    int retired = 0;            // items whose stores have been issued ...
    int rt_s = 0, rt_ph = 0;    // ... slot and mbarrier phase of the next one to retire
    // retire the oldest unretired item: wait for the consumers, then write the stage back
    auto retire = [&]() {
      mbar_wait(&s_done[rt_s], (uint32_t)rt_ph);
      const BulkItem& b = s_item[rt_s];
      if (b.nbulk > 0) {
        float* sp = reinterpret_cast<float*>(s_ring + rt_s * stage_bytes);
        const uint32_t bytes = (uint32_t)b.nbulk * 4u;
        bulk_store(b.P, sp, bytes);
        if (zero_grads) bulk_store(b.G, sp + tile, bytes);
        if (b.M) bulk_store(b.M, sp + 2 * tile, bytes);
        if (Op::kHasV) bulk_store(b.V, sp + 3 * tile, bytes);
      }
      bulk_commit();           // one group per item, empty or not: keeps the accounting uniform
      ++retired;
      if (++rt_s == stages) { rt_s = 0; rt_ph ^= 1; }
    };
    auto enqueue = [&](const BulkItem& b) {
      // the slot of item `it` was last used by item it - stages: its stores must have been
      // issued (retire) and must have finished reading shared memory
      if (it >= stages) {
        while (retired <= it - stages) retire();
        // groups committed after the one of item (it - stages) may stay pending
        const int younger = retired - 1 - (it - stages);
        if (younger <= 0) bulk_wait_read<0>();
        else if (younger == 1) bulk_wait_read<1>();
        else if (younger == 2) bulk_wait_read<2>();
        else bulk_wait_read<3>();     // conservative for deeper lags
      }
      s_item[it_s] = b;
      if (b.nbulk > 0) {
        float* sp = reinterpret_cast<float*>(s_ring + it_s * stage_bytes);
        const uint32_t bytes = (uint32_t)b.nbulk * 4u;
        const int n_streams = 2 + (b.M != nullptr) + (Op::kHasV ? 1 : 0);
        mbar_arrive_expect_tx(&s_full[it_s], bytes * n_streams);
        bulk_load(sp, b.P, bytes, &s_full[it_s]);
        bulk_load(sp + tile, b.G, bytes, &s_full[it_s]);
        if (b.M) bulk_load(sp + 2 * tile, b.M, bytes, &s_full[it_s]);
        if (Op::kHasV) bulk_load(sp + 3 * tile, b.V, bytes, &s_full[it_s]);
      } else {
        mbar_arrive(&s_full[it_s]);
      }
      ++it;
      if (++it_s == stages) it_s = 0;
    };
    const int lag = stages > 2 ? stages - 2 : 1;   // stages kept loading ahead of the stores
    // the tensor of the previous item is usually the tensor of this one: keep its descriptor
    int t_lo = 0, t_hi = 0;            // chunk range [t_lo, t_hi) of the cached tensor
    rk_tensor_desc d;
    bool d_skip = true;
    for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
      const int c = w >> sub_log2;
      const int sub = w & sub_mask;
      if (c < t_lo || c >= t_hi) {
        const int t = rk_find_tensor(prefix, n_tensors, c);
        t_lo = prefix[t];
        t_hi = prefix[t + 1];
        d = descs[t];
        const int gflags = dtable ? dtable->g[d.group].flags : table.g[d.group].flags;
        d_skip = (d.flags & RK_TENSOR_SKIP) || (gflags & GROUP_SKIP);
      }
      if (d_skip) continue;
      const int64_t base = (int64_t)(c - t_lo) * chunk + ((int64_t)sub << tile_log2);
      const int64_t rem = d.numel - base;
      if (rem <= 0) continue;
      BulkItem b;
      b.count = rem < tile ? (int)rem : tile;
      b.P = d.p + base;
      b.G = d.g + base;
      b.M = d.m ? d.m + base : nullptr;
      b.V = Op::kHasV ? d.v + base : nullptr;
      b.group = d.group;
      b.mode = 0;
      b.nbulk = (d.flags & RK_TENSOR_VEC4) ? (b.count & ~3) : 0;
      // write back finished tiles early enough that `lag` loads stay in flight
      while (retired < it - lag) retire();
      enqueue(b);
    }
    BulkItem e;
    e.P = e.G = e.M = e.V = nullptr;
    e.count = e.nbulk = e.group = 0;
    e.mode = 1;
    enqueue(e);
    while (retired < it - 1) retire();   // everything before the end marker
    bulk_wait_all();
    return;
  }

  // ================= consumers =================
  grid_dependency_wait();      // the norm below is the previous kernel's output
  Prolog pr;
  pr.scale = grad_scale;
  pr.has_scale = grad_scale != 1.0f;
  pr.has_clip = total_norm != nullptr;
  pr.coef = pr.has_clip ? rk_clip_coef(__ldcg(total_norm), max_norm) : 1.0f;
  const int ctid = tid - 32;
  constexpr int NC = CW * 32;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  const int tile4 = tile >> 2;
  int s = 0;
  uint32_t ph = 0;
  for (;;) {
    mbar_wait(&s_full[s], ph);
    const BulkItem b = s_item[s];
    if (b.mode == 1) break;
    Op op;
    op.h = dtable ? dtable->g[b.group] : table.g[b.group];
    const bool has_m = b.M != nullptr;
    float4* sp = reinterpret_cast<float4*>(s_ring + s * stage_bytes);
    float4* sg = sp + tile4;
    float4* sm = sg + tile4;
    float4* sv = sm + tile4;
    const int nvec = b.nbulk >> 2;
    for (int i = ctid; i < nvec; i += NC) {
      float4 p4 = sp[i], g4 = sg[i], m4 = zero4, v4 = zero4;
      if (has_m) m4 = sm[i];
      if (Op::kHasV) v4 = sv[i];
      op(p4.x, prolog_g(g4.x, pr), m4.x, v4.x);
      op(p4.y, prolog_g(g4.y, pr), m4.y, v4.y);
      op(p4.z, prolog_g(g4.z, pr), m4.z, v4.z);
      op(p4.w, prolog_g(g4.w, pr), m4.w, v4.w);
      sp[i] = p4;
      if (has_m) sm[i] = m4;
      if (Op::kHasV) sv[i] = v4;
      if (zero_grads) sg[i] = zero4;
    }
    // elements that did not go through shared memory: the < 4 tail, or a whole unaligned tile
    for (int i = b.nbulk + ctid; i < b.count; i += NC) {
      float p = ldg1(b.P + i), m = has_m ? ldg1(b.M + i) : 0.f, v = Op::kHasV ? ldg1(b.V + i) : 0.f;
      op(p, prolog_g(ldg1(b.G + i), pr), m, v);
      stg1(b.P + i, p);
      if (has_m) stg1(b.M + i, m);
      if (Op::kHasV) stg1(b.V + i, v);
      if (zero_grads) stg1(b.G + i, 0.f);
    }
    fence_async_smem();        // my shared-memory writes -> visible to the bulk stores
    __syncwarp();
    if (lane == 0) mbar_arrive(&s_done[s]);
    if (++s == stages) { s = 0; ph ^= 1; }
  }
}
