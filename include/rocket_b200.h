/*
 * rocket_b200.h -- C-ABI of librocket_b200.so
 *
 * Hand-written sm_100a (B200) kernels for the engine-owned part of rocket's
 * per-iteration training step.  The reference (dsenushkin/rocket) is pure
 * Python and has no FFI of its own: every entry point below replaces a call
 * the reference makes into torch / accelerate on its hot path, and cites that
 * call site (paths relative to the reference tree, $TORCH = site-packages/torch).
 *
 * Conventions (all entry points):
 *   - plain C types only; no torch types; `stream` is a cudaStream_t passed as void*.
 *   - every pointer is caller-owned.  The library never allocates or frees device
 *     memory.  Device pointers must belong to the device that is current
 *     when the call is made.
 *   - every call that takes a stream is asynchronous on it and never synchronises
 *     the host (exception: rk_stage_wait blocks the HOST until the copy has been
 *     *issued*, not until it has finished).
 *   - return value: 0 on success, a negative rk_status on a usage error, a positive
 *     cudaError_t on a CUDA error.  rk_last_error() gives a thread-local message.
 *   - nothing here throws.
 */
#ifndef ROCKET_B200_H_
#define ROCKET_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RK_VERSION 100 /* 0.1.0 */

typedef enum rk_status {
  RK_OK = 0,
  RK_ERR_INVALID = -1,  /* bad argument (null pointer, negative size, ...) */
  RK_ERR_TOO_LARGE = -2,/* a size exceeds a documented limit                */
  RK_ERR_STATE = -3,    /* object used in the wrong state                   */
  RK_ERR_UNSUPPORTED = -4
} rk_status;

/* ---- limits ---------------------------------------------------------------- */
#define RK_MAX_GROUPS 64          /* optimizer param groups per launch           */
#define RK_STEP_CHUNK 4096        /* elements per work chunk of the step kernels */
#define RK_STEP_CHUNK_SMALL 1024  /* ... for tensor lists under 2^26 elements   */
#define RK_NORM_CHUNK_MIN 4096    /* smallest chunk of the norm kernel           */

/* ---- tensor list ----------------------------------------------------------- */
/* One fp32 parameter tensor and its optimizer state.  `m` is exp_avg (Adam) or
 * momentum_buffer (SGD); `v` is exp_avg_sq (Adam) and unused for SGD.  Replaces the
 * per-parameter Python lists torch builds in Optimizer._init_group
 * ($TORCH/optim/adam.py:134-205).                                               */
typedef struct rk_tensor_desc {
  float* p;
  float* g;
  float* m;
  float* v;
  int64_t numel;
  int32_t group; /* index into the hyper-parameter table, < RK_MAX_GROUPS       */
  int32_t flags; /* RK_TENSOR_SKIP: leave this tensor untouched (grad is None)  */
} rk_tensor_desc;
#define RK_TENSOR_SKIP 1

/* Adam / AdamW hyper-parameters of one param group, exactly the values
 * torch.optim.Adam(W) keeps in `param_groups` ($TORCH/optim/adam.py:34-90) plus the
 * step count torch keeps per parameter (adam.py:168-175).  `step` is the count AFTER
 * the increment at adam.py:414, i.e. 1 for the first update.  The library derives
 * bias corrections and step size in double precision on the host the way
 * _single_tensor_adam does (adam.py:529-536).                                   */
typedef struct rk_adam_group {
  double lr, beta1, beta2, eps, weight_decay;
  int64_t step;
  int32_t decoupled; /* 1 = AdamW (adam.py:419), 0 = Adam with L2 (adam.py:430) */
  int32_t skip;      /* 1 = do not update tensors of this group                 */
} rk_adam_group;

/* SGD hyper-parameters of one param group ($TORCH/optim/sgd.py:27-70).
 * first_step=1 means momentum buffers do not exist yet: buf = grad (sgd.py:362).*/
typedef struct rk_sgd_group {
  double lr, momentum, dampening, weight_decay;
  int32_t nesterov;
  int32_t first_step;
  int32_t skip;
  int32_t _pad;
} rk_sgd_group;

/* ---- library ---------------------------------------------------------------- */
int rk_version(void);
const char* rk_last_error(void);
/* Number of kernels this library has launched in this process (for bench.py's
 * gpu_launches claim).                                                          */
uint64_t rk_launch_count(void);

/* ---- plan: tensor-list metadata ---------------------------------------------
 * A plan is a position-independent blob: header, descriptor table and the chunk
 * prefix tables the kernels binary-search (staged into shared memory per CTA).
 * Build it once on the host, copy the same bytes to device memory, and pass both
 * copies to the kernels (host copy sizes the grid; device copy is read by the GPU).
 * scratch_bytes is the device scratch rk_l2norm_multi needs; it must be zero-filled
 * once before its first use.
 * Replaces: the tensor-list grouping torch redoes every step in
 * _group_tensors_by_device_and_dtype ($TORCH/nn/utils/clip_grad.py:85-90).     */
int rk_plan_query(const int64_t* numel, int n_tensors, size_t* plan_bytes,
                  size_t* scratch_bytes);
int rk_plan_build(const rk_tensor_desc* descs, int n_tensors, void* plan_host,
                  size_t plan_bytes);

/* ---- K1: global gradient L2 norm ---------------------------------------------
 * total = || ( ||g_0 * s||_2, ||g_1 * s||_2, ... ) ||_2  with s = grad_scale, the
 * two-level form of torch.nn.utils.clip_grad_norm_ / _get_total_norm
 * ($TORCH/nn/utils/clip_grad.py:96-107), which accelerate's clip_grad_norm_ calls.
 * Deterministic: per-chunk partials in a fixed order, no float atomics, so every
 * data-parallel rank derives a bit-identical clip coefficient.  Two launches on `stream`:
 * level 1 (partials) and the one-CTA level 2 of rk_l2norm_finalize.
 * Writes total to *total_norm_dev (device float) and, if non-null, per-tensor norms
 * to tensor_norms_dev[n_tensors].  Tensors flagged RK_TENSOR_SKIP contribute 0.   */
int rk_l2norm_multi(const void* plan_host, const void* plan_dev, float grad_scale,
                    void* scratch_dev, float* total_norm_dev,
                    float* tensor_norms_dev, void* stream);

/* ---- K1 again, for callers that cannot supply scratch memory -------------------------------
 * No counterpart in the reference: this entry point exists only because of who may call
 * this library.  Some hosts cannot allocate device memory at the point where they need the
 * norm: a runtime that handed the whole GPU to its own memory pool at start-up and forbids
 * driver allocations afterwards, or a thin foreign-language binding (cgo, JNI, a plugin
 * ABI) that only ever holds the device pointers its framework passes in -- the gradients
 * and a place for the result -- and has no way to create or own a buffer of its own.
 *
 * Same kernel and same result as rk_l2norm_multi, but the scratch (per-chunk partials,
 * per-tensor norms, completion ticket) is a fixed array inside the library's device image:
 * RK_STATIC_SCRATCH_BYTES on every GPU that loads the library, held until the process
 * exits.  The call allocates nothing on the device, now or later (no cudaMalloc, no
 * cudaFree, no allocating library routine); the caller still passes the plan, the
 * gradients it describes and the output pointers.
 *   - Tensor lists whose plan needs more scratch than that (rk_plan_query's
 *     scratch_bytes; roughly 8 bytes per norm chunk + 4 per tensor, i.e. beyond ~100K
 *     tensors) are refused with RK_ERR_TOO_LARGE and a message; nothing is launched.
 *   - Every call looks the array's device address up afresh (the device the call is for must
 *     be current, and `stream` must belong to it: neither is checked), so the entry point
 *     keeps working after cudaDeviceReset or a re-created primary context.  The first call on
 *     each device additionally does one-time set-up (it loads the kernel and sizes its grid):
 *     make that first call before starting any stream capture.
 *   - All calls on a device share the one array.  It keeps the previous call's contents;
 *     the entry point clears the part that must start at zero itself.  Two calls whose
 *     kernels can overlap on the GPU (different streams, or different host threads without
 *     ordering between them) race on it: serialise them on one stream or with events.
 *   - total_norm_dev must not be NULL (there is no caller-visible scratch to read the
 *     result from).                                                                        */
#define RK_STATIC_SCRATCH_BYTES (1u << 20)
int rk_l2norm_multi_static(const void* plan_host, const void* plan_dev, float grad_scale,
                           float* total_norm_dev, float* tensor_norms_dev, void* stream);

/* Level 2 of K1 on its own: per-tensor norms and the total from per-chunk partials that are
 * already in scratch (rk_comm_allreduce_norm writes them).  Same fixed-order reduction: it is
 * the very kernel rk_l2norm_multi launches behind its level-1 kernel (one CTA, launched as a
 * programmatic dependent so that its set-up overlaps the producer's tail), hence the same
 * bits whichever kernel produced the partials.                                            */
int rk_l2norm_finalize(const void* plan_host, const void* plan_dev, void* scratch_dev,
                       float* total_norm_dev, float* tensor_norms_dev, void* stream);

/* ---- K2: fused multi-tensor Adam / AdamW step ---------------------------------
 * One pass over HBM doing, per element and in this order:
 *   g  = g * grad_scale                      (1/world_size; skipped when 1.0)
 *   g  = g * clip_coef                       clip_grad.py:165-182; only if total_norm_dev
 *   p *= 1 - lr*wd   |  g += wd*p            adam.py:419 | adam.py:430
 *   m  = lerp(m, g, 1-beta1)                 adam.py:457
 *   v  = v*beta2 + (1-beta2)*g*g             adam.py:476
 *   p += -step_size * m / (sqrt(v)/sqrt(bc2) + eps)   adam.py:545-547
 *   g  = 0                                   if zero_grads (Optimizer.zero_grad,
 *                                            rocket/core/optimizer.py:130)
 * clip_coef = min(max_norm * 1/(total_norm + 1e-6), 1) is derived on the device from
 * *total_norm_dev (no host sync).  fp32 rounding sequence follows torch's CPU kernels
 * (foreach=False) bit for bit; see oracle/step_oracle.c.
 * Replaces rocket/core/optimizer.py:129-130 -> torch.optim.AdamW.step + zero_grad. */
int rk_adamw_step_multi(const void* plan_host, const void* plan_dev,
                        const rk_adam_group* groups, int n_groups,
                        const float* total_norm_dev, float max_norm,
                        float grad_scale, int zero_grads, void* stream);

/* The nine fp32 constants K2 uses for one group, derived in double as
 * _single_tensor_adam does (adam.py:419,529-536): {1-lr*wd, wd, 1-beta1, (1-beta1)-1,
 * beta2, 1-beta2, sqrt(bias_correction2), eps, -lr/bias_correction1}.  Host only.   */
int rk_adam_derive(const rk_adam_group* group, float* out9, int* flags_out);

/* ---- K3: fused multi-tensor SGD step -------------------------------------------
 * Same extras (grad_scale, clip, zero_grads) around _single_tensor_sgd
 * ($TORCH/optim/sgd.py:343-380): weight decay, momentum/dampening, nesterov.     */
int rk_sgd_step_multi(const void* plan_host, const void* plan_dev,
                      const rk_sgd_group* groups, int n_groups,
                      const float* total_norm_dev, float max_norm,
                      float grad_scale, int zero_grads, void* stream);

/* Tuning hook (no counterpart in the reference): K2/K3 exist in several code shapes --
 * per-thread 128-bit or 256-bit ld/st.global walkers at different occupancies, and a
 * bulk-copy (cp.async.bulk + mbarrier ring) kernel -- all running the same element code,
 * hence the same bits.  The library picks one by list size; `variant >= 0` forces one
 * process-wide (the numbers are listed beside the switch in csrc/rk_step.cu), `-1` restores
 * the automatic choice.  Environment RK_STEP_VARIANT sets the initial value.             */
int rk_step_variant_override(int variant);

/* ---- K2/K3 with the per-group constants in device memory ---------------------------------
 * For CUDA-graph replays, where kernel arguments are frozen at capture: rk_*_table_write
 * derives the same per-group constants as the calls above into RK_STEP_TABLE_BYTES of
 * caller memory (host side; typically a pinned slot), the caller copies them to
 * `table_dev` in stream order, and the *_devtable launches read them from there.  Same
 * kernels, same arithmetic.                                                              */
#define RK_STEP_TABLE_BYTES 4096
int rk_adam_table_write(const rk_adam_group* groups, int n_groups, void* table_out);
int rk_sgd_table_write(const rk_sgd_group* groups, int n_groups, void* table_out);
int rk_adamw_step_multi_devtable(const void* plan_host, const void* plan_dev,
                                 const void* table_dev, const float* total_norm_dev,
                                 float max_norm, float grad_scale, int zero_grads, void* stream);
int rk_sgd_step_multi_devtable(const void* plan_host, const void* plan_dev,
                               const void* table_dev, const float* total_norm_dev,
                               float max_norm, float grad_scale, int zero_grads, void* stream);

/* ---- K4: Meter reductions on the device -----------------------------------------
 * counters[0] += #rows whose argmax(logits[row, :]) == labels[row];
 * counters[1] += n_rows.   Integer atomics: bit-exact and order independent.
 * argmax follows torch.argmax: first index wins ties, NaN counts as the maximum.
 * dtype: 0 = fp32, 1 = bf16, 2 = fp16.  row_stride in elements.
 * Replaces the per-step `argmax(1) / == / sum / .item()` of the example metric
 * (examples/mnist.py:26-34) -- no host sync.                                      */
int rk_meter_argmax_count(const void* logits, int dtype, const int64_t* labels,
                          int64_t n_rows, int64_t n_cols, int64_t row_stride,
                          long long* counters_dev, void* stream);
/* *acc_dev += (double)(*value_dev) / divisor  -- the loss bookkeeping of
 * rocket/core/loss.py:95-98 (`value += loss.item() / accumulation_steps`) without
 * the .item() sync.                                                               */
int rk_scalar_accumulate(const float* value_dev, double divisor, double* acc_dev,
                         void* stream);

/* ---- K6: gradient all-reduce over NVLink peer memory, fused with 1/W and the norm ------------
 * Replaces what the reference gets from torch DDP (rocket/core/module.py:106 ->
 * accelerator.prepare -> DistributedDataParallel: $TORCH/nn/parallel/distributed.py, C++
 * reducer + ncclAllReduce per 25 MiB bucket) and the separate norm pass of clip_grad_norm_.
 * Every rank owns one device "arena" that all ranks of the node have mapped (CUDA IPC; the
 * host side does the mapping):   [ signal pad: RK_COMM_SIGNAL_BYTES | norm scratch | gradients ]
 * arenas[q] is rank q's arena as mapped in THIS process (arenas[rank] is the local one);
 * scratch_offset is where the norm scratch starts inside every arena.  The gradient pointers
 * in the plan must point into the local arena; the same offsets are used on the peers.
 * rk_comm_allreduce_norm reduces the norm chunks [chunk_lo, chunk_hi) of the plan (a bucket):
 * afterwards every rank holds sum_r(g_r) * (1/W) in place, bit-identical everywhere, and the
 * chunks' sum-of-squares partials in its scratch; rk_l2norm_finalize turns them into the
 * norm.  All ranks must call it with the same arguments in the same order; each rank's
 * kernel waits for its peers' (flags with a per-bucket epoch, no resets), so every rank's
 * launch must be able to run concurrently with the others' -- one GPU per rank.
 * world in {2, 4, 8}; bucket < RK_COMM_MAX_BUCKETS; max_ctas <= 0 means one CTA per SM.
 * multicast_arena (may be NULL): an NVSwitch multicast mapping of all W arenas.  With it and
 * use_multicast != 0 the kernel reduces with multimem.ld_reduce and replicates with
 * multimem.st (in-switch sum / broadcast) instead of W loads and W stores per element.      */
#define RK_COMM_MAX_WORLD 8
#define RK_COMM_MAX_BUCKETS 256
#define RK_COMM_SIGNAL_BYTES 36864
typedef struct rk_comm rk_comm;
int rk_comm_create(rk_comm** out, int rank, int world, void* const* arenas,
                   void* multicast_arena, uint64_t scratch_offset);
int rk_comm_allreduce_norm(rk_comm* comm, const void* plan_host, const void* plan_dev,
                           int chunk_lo, int chunk_hi, int bucket, int max_ctas,
                           int use_multicast, void* stream);
int rk_comm_destroy(rk_comm* comm);

/* ---- K7: sharded optimizer step -----------------------------------
 * The reference has no counterpart: under DDP every replica repeats the whole
 * torch.optim step (rocket/core/optimizer.py:129).  Here the arena additionally carries the
 * parameters   [ signal pad | norm scratch | gradients | parameters ]   (the plan's p pointers,
 * like its g pointers, must point into the local arena; the same offsets are used on the
 * peers) and rank r owns the norm chunks c with c % W == r, the same ownership as the exchange.
 * rk_comm_reduce_scatter_norm is rk_comm_allreduce_norm without the broadcast half: the owner
 * alone keeps sum_r(g_r) * (1/W) of its chunks (other ranks' copies of those chunks stay as
 * they were); the norm partials still reach every rank.
 * rk_comm_sharded_step then applies rk_adamw_step_multi_devtable's (optimizer 0) or
 * rk_sgd_step_multi_devtable's (optimizer 1) element update to the owned chunks only, writes
 * the new parameters into every rank's arena (P2P stores / multimem.st), keeps m and v local,
 * zero-fills the whole local gradient buffer, and ends with a barrier among the ranks.  Per
 * element the arithmetic is the replicated step's, so parameters are bit-identical to data
 * parallelism; optimizer work and moment traffic per GPU shrink by W.  table_dev is the
 * RK_STEP_TABLE_BYTES table written by rk_adam_table_write / rk_sgd_table_write.
 * Moments: with shard_m / shard_v == NULL the plan's per-tensor m / v buffers are used (full
 * size; the parts of chunks a rank does not own are never touched).  Otherwise they are this
 * rank's compact moment shards, 1/W of the full size: the moments of owned norm chunk c occupy
 * elements [(c / W) * chunk_norm, ... + chunk length) -- ceil(n_chunks_norm / W) * chunk_norm
 * floats each, 16-byte aligned, zero-filled before the first step -- and the plan's m / v
 * pointers are ignored (shard_v is ignored for SGD; shard_m == NULL with plan m == NULL means
 * plain SGD without momentum).  grad_scale multiplies every gradient as it is read (1.0, or
 * 1/loss_scale under fp16: the exchange has already averaged).  Uses barrier slot RK_COMM_MAX_BUCKETS - 1: exchanges that
 * share the communicator must use lower buckets.                                            */
int rk_comm_reduce_scatter_norm(rk_comm* comm, const void* plan_host, const void* plan_dev,
                                int chunk_lo, int chunk_hi, int bucket, int max_ctas,
                                int use_multicast, void* stream);
int rk_comm_sharded_step(rk_comm* comm, const void* plan_host, const void* plan_dev,
                         int optimizer, const void* table_dev, float* shard_m, float* shard_v,
                         const float* total_norm_dev, float max_norm, float grad_scale,
                         int max_ctas, int use_multicast, void* stream);

/* ---- K5: pinned-ring batch stager ---------------------------------------------------
 * Replaces the blocking pageable `Tensor.to(device)` of rocket/utils/torch.py:50-51
 * reached from rocket/core/dataset.py:279-282.  The caller supplies a pinned host ring
 * and a device ring of n_slots * slot_bytes each; the stager owns a copy stream, events
 * and a worker thread that packs host segments into the pinned slot and issues ONE
 * cudaMemcpyAsync per batch.
 *   submit : enqueue; returns at once.  src pointers must stay valid until wait returns.
 *   wait   : host blocks until the copy is issued; `consumer` stream then waits for it.
 *   release: records on `consumer` that the slot's device memory may be overwritten.   */
typedef struct rk_stage rk_stage;
int rk_stage_create(rk_stage** out, void* pinned_base, void* device_base,
                    size_t slot_bytes, int n_slots, int n_copy_threads);
int rk_stage_submit(rk_stage* st, int slot, const void* const* src,
                    const size_t* offsets, const size_t* sizes, int n_segments,
                    uint64_t* ticket);
int rk_stage_wait(rk_stage* st, uint64_t ticket, void* consumer_stream);
int rk_stage_release(rk_stage* st, int slot, void* consumer_stream);
int rk_stage_destroy(rk_stage* st);

#ifdef __cplusplus
}
#endif
#endif /* ROCKET_B200_H_ */
