// Library-level entry points: version, errors, launch counter, plan construction.
#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

#include "rk_common.cuh"

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_rk_launches{0};

void rk_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

extern "C" int rk_version(void) { return RK_VERSION; }
extern "C" const char* rk_last_error(void) { return t_err; }
extern "C" uint64_t rk_launch_count(void) { return g_rk_launches.load(); }

// ---- persistent grid sizing ---------------------------------------------------
// Cached per (device, kernel): resident CTAs per SM x SM count.
int rk_persistent_grid(const void* kernel, int block, size_t dyn_smem, int n_work,
                       int* grid_out) {
  struct Key {
    int dev;
    const void* k;
    bool operator==(const Key& o) const { return dev == o.dev && k == o.k; }
  };
  struct Hash {
    size_t operator()(const Key& k) const {
      return std::hash<const void*>()(k.k) ^ (size_t)k.dev * 0x9e3779b97f4a7c15ull;
    }
  };
  static std::mutex mu;
  static std::unordered_map<Key, int, Hash> cache;
  int dev = 0;
  RK_CUDA(cudaGetDevice(&dev));
  int resident = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(Key{dev, kernel});
    if (it != cache.end()) resident = it->second;
  }
  if (resident == 0) {
    int sms = 0, per_sm = 0;
    RK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    RK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, dyn_smem));
    RK_REQUIRE(per_sm > 0 && sms > 0, RK_ERR_STATE, "kernel does not fit on an SM");
    resident = sms * per_sm;
    std::lock_guard<std::mutex> lk(mu);
    cache[Key{dev, kernel}] = resident;
  }
  *grid_out = n_work < resident ? (n_work > 0 ? n_work : 1) : resident;
  return 0;
}

// ---- plan ---------------------------------------------------------------------
static int norm_chunk_for(int64_t total) {
  // Aim for <= ~16K norm chunks: level 1 runs one block-wide reduction per chunk, so a chunk
  // should be at least two load batches (8192 elements) once the list is large, but
  // keep chunks small enough that 148 SMs x resident CTAs balance.
  int64_t c = RK_NORM_CHUNK_MIN;
  while (total / c > 16384 && c < (1 << 20)) c <<= 1;
  return (int)c;
}

struct PlanLayout {
  rk_plan_header h;
};

static int layout_plan(const int64_t* numel, int n, PlanLayout* out) {
  RK_REQUIRE(n > 0, RK_ERR_INVALID, "rk_plan: n_tensors must be > 0 (got %d)", n);
  RK_REQUIRE(numel != nullptr, RK_ERR_INVALID, "rk_plan: numel is null");
  int64_t total = 0;
  for (int i = 0; i < n; ++i) {
    RK_REQUIRE(numel[i] >= 0, RK_ERR_INVALID, "rk_plan: tensor %d has negative numel", i);
    total += numel[i];
  }
  rk_plan_header& h = out->h;
  memset(&h, 0, sizeof(h));
  h.magic = RK_PLAN_MAGIC;
  h.n_tensors = n;
  // small lists get finer chunks: with ~444 resident CTAs a 11M-parameter model is only ~6
  // waves of 4096-element chunks, i.e. a 12 % tail; 1024-element chunks cut that to < 2 %
  h.chunk_step = total >= ((int64_t)1 << 26) ? RK_STEP_CHUNK : RK_STEP_CHUNK_SMALL;
  h.chunk_norm = norm_chunk_for(total);
  h.total_numel = total;
  int64_t cs = 0, cn = 0;
  for (int i = 0; i < n; ++i) {
    cs += (numel[i] + h.chunk_step - 1) / h.chunk_step;
    cn += (numel[i] + h.chunk_norm - 1) / h.chunk_norm;
  }
  RK_REQUIRE(cs < (int64_t)INT32_MAX, RK_ERR_TOO_LARGE,
             "rk_plan: %lld step chunks exceed int32", (long long)cs);
  h.n_chunks_step = (int)cs;
  h.n_chunks_norm = (int)cn;
  size_t off = rk_align_up(sizeof(rk_plan_header), 16);
  h.off_descs = (uint32_t)off;
  off = rk_align_up(off + sizeof(rk_tensor_desc) * (size_t)n, 16);
  h.off_prefix_step = (uint32_t)off;
  off = rk_align_up(off + sizeof(int32_t) * (size_t)(n + 1), 16);
  h.off_prefix_norm = (uint32_t)off;
  off = rk_align_up(off + sizeof(int32_t) * (size_t)(n + 1), 16);
  RK_REQUIRE(off < UINT32_MAX, RK_ERR_TOO_LARGE, "rk_plan: plan blob too large");
  h.plan_bytes = (uint32_t)off;
  size_t so = 0;
  h.off_partials = 0;
  so = rk_align_up(sizeof(double) * (size_t)(cn > 0 ? cn : 1), 256);
  h.off_tnorms = (uint32_t)so;
  so = rk_align_up(so + sizeof(float) * (size_t)n, 256);
  h.off_total = (uint32_t)so;
  so += 128;
  h.off_ticket = (uint32_t)so;
  so += 128;
  h.scratch_bytes = (uint32_t)so;
  return 0;
}

extern "C" int rk_plan_query(const int64_t* numel, int n_tensors, size_t* plan_bytes,
                             size_t* scratch_bytes) {
  PlanLayout L;
  int rc = layout_plan(numel, n_tensors, &L);
  if (rc) return rc;
  if (plan_bytes) *plan_bytes = L.h.plan_bytes;
  if (scratch_bytes) *scratch_bytes = L.h.scratch_bytes;
  return 0;
}

extern "C" int rk_plan_build(const rk_tensor_desc* descs, int n_tensors, void* plan_host,
                             size_t plan_bytes) {
  RK_REQUIRE(descs && plan_host, RK_ERR_INVALID, "rk_plan_build: null argument");
  RK_REQUIRE(n_tensors > 0, RK_ERR_INVALID, "rk_plan_build: n_tensors must be > 0");
  int64_t* numel = new int64_t[n_tensors];
  for (int i = 0; i < n_tensors; ++i) numel[i] = descs[i].numel;
  PlanLayout L;
  int rc = layout_plan(numel, n_tensors, &L);
  delete[] numel;
  if (rc) return rc;
  RK_REQUIRE(plan_bytes >= L.h.plan_bytes, RK_ERR_INVALID,
             "rk_plan_build: buffer of %zu bytes, need %u", plan_bytes, L.h.plan_bytes);
  char* base = (char*)plan_host;
  memset(base, 0, L.h.plan_bytes);
  memcpy(base, &L.h, sizeof(L.h));
  rk_tensor_desc* d = (rk_tensor_desc*)(base + L.h.off_descs);
  int32_t* ps = (int32_t*)(base + L.h.off_prefix_step);
  int32_t* pn = (int32_t*)(base + L.h.off_prefix_norm);
  int32_t cs = 0, cn = 0;
  for (int i = 0; i < n_tensors; ++i) {
    const rk_tensor_desc& s = descs[i];
    RK_REQUIRE(s.group >= 0 && s.group < RK_MAX_GROUPS, RK_ERR_INVALID,
               "rk_plan_build: tensor %d group %d out of range [0,%d)", i, s.group,
               RK_MAX_GROUPS);
    RK_REQUIRE(s.numel == 0 || (s.flags & RK_TENSOR_SKIP) || (s.p && s.g), RK_ERR_INVALID,
               "rk_plan_build: tensor %d has a null p or g pointer", i);
    d[i] = s;
    d[i].flags &= RK_TENSOR_SKIP;
    uintptr_t bits = (uintptr_t)s.p | (uintptr_t)s.g | (uintptr_t)s.m | (uintptr_t)s.v;
    if ((bits & 15) == 0) d[i].flags |= RK_TENSOR_VEC4;
    if ((bits & 31) == 0) d[i].flags |= RK_TENSOR_VEC8;
    ps[i] = cs;
    pn[i] = cn;
    cs += (int32_t)((s.numel + L.h.chunk_step - 1) / L.h.chunk_step);
    cn += (int32_t)((s.numel + L.h.chunk_norm - 1) / L.h.chunk_norm);
  }
  ps[n_tensors] = cs;
  pn[n_tensors] = cn;
  return 0;
}
