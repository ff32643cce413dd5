// K5: pinned-ring batch stager.
//
// A worker thread (no GIL) packs the host segments of one batch into a pinned slot and
// issues ONE cudaMemcpyAsync per batch on a dedicated copy stream; events order the copy
// against the consumer (compute) stream in both directions:
//     copied[slot]   : recorded on the copy stream after the H2D; the consumer waits on it.
//     released[slot] : recorded on the consumer stream when the batch is dropped; the copy
//                      stream waits on it before the slot's device memory is overwritten.
// So batch i+1 is packed and copied while batch i is being computed on, and nothing on the
// host ever waits for a copy to finish -- only for it to be issued.
#include <string.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "rk_common.cuh"

namespace {

struct Segment {
  const char* src;
  size_t off, size;
};
struct Job {
  int slot;
  uint64_t ticket;
  std::vector<Segment> segs;
  size_t extent;
};

void pack_range(const std::vector<Segment>& segs, char* dst, size_t lo, size_t hi) {
  // copy bytes [lo, hi) of the concatenation of all segments
  size_t pos = 0;
  for (const Segment& s : segs) {
    const size_t a = pos, b = pos + s.size;
    pos = b;
    if (b <= lo) continue;
    if (a >= hi) break;
    const size_t from = lo > a ? lo - a : 0;
    const size_t to = (hi < b ? hi : b) - a;
    memcpy(dst + s.off + from, s.src + from, to - from);
  }
}

}  // namespace

struct rk_stage {
  char* pinned = nullptr;
  char* dev = nullptr;
  size_t slot_bytes = 0;
  int n_slots = 0, n_threads = 1, device = 0;
  cudaStream_t copy_stream = nullptr;
  std::vector<cudaEvent_t> copied, released;
  std::vector<char> copied_valid, released_valid;
  uint64_t ticket_slot[64];
  std::thread worker;
  std::mutex mu;
  std::condition_variable cv_job, cv_done;
  std::deque<Job> q;
  uint64_t next_ticket = 1, issued = 0;
  bool stop = false;
  int err = 0;
  std::string err_msg;

  void fail(int code, const char* what) {
    std::lock_guard<std::mutex> lk(mu);
    if (!err) {
      err = code;
      err_msg = what;
    }
  }

  void run() {
    if (cudaSetDevice(device) != cudaSuccess) fail(RK_ERR_STATE, "stager: cudaSetDevice failed");
    for (;;) {
      Job job;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv_job.wait(lk, [&] { return stop || !q.empty(); });
        if (q.empty()) return;
        job = std::move(q.front());
        q.pop_front();
      }
      char* hdst = pinned + (size_t)job.slot * slot_bytes;
      char* ddst = dev + (size_t)job.slot * slot_bytes;
      cudaError_t e = cudaSuccess;
      bool was_copied, was_released;
      {
        std::lock_guard<std::mutex> lk(mu);
        was_copied = copied_valid[job.slot];
        was_released = released_valid[job.slot];
      }
      // the previous H2D out of this pinned slot must have drained before we overwrite it
      if (was_copied) e = cudaEventSynchronize(copied[job.slot]);
      size_t total = 0;
      for (const Segment& s : job.segs) total += s.size;
      const int nt = (total >= (size_t)8 << 20) ? n_threads : 1;
      if (nt <= 1) {
        pack_range(job.segs, hdst, 0, total);
      } else {
        std::vector<std::thread> ts;
        const size_t per = rk_align_up((total + nt - 1) / nt, 4096);
        for (int i = 1; i < nt; ++i) {
          const size_t lo = per * i, hi = lo + per < total ? lo + per : total;
          if (lo < total) ts.emplace_back(pack_range, std::cref(job.segs), hdst, lo, hi);
        }
        pack_range(job.segs, hdst, 0, per < total ? per : total);
        for (auto& t : ts) t.join();
      }
      if (e == cudaSuccess && was_released)
        e = cudaStreamWaitEvent(copy_stream, released[job.slot], 0);
      if (e == cudaSuccess && job.extent)
        e = cudaMemcpyAsync(ddst, hdst, job.extent, cudaMemcpyHostToDevice, copy_stream);
      if (e == cudaSuccess) e = cudaEventRecord(copied[job.slot], copy_stream);
      if (e != cudaSuccess) fail((int)e, cudaGetErrorString(e));
      {
        std::lock_guard<std::mutex> lk(mu);
        copied_valid[job.slot] = 1;
        issued = job.ticket;
      }
      cv_done.notify_all();
    }
  }
};

extern "C" int rk_stage_create(rk_stage** out, void* pinned_base, void* device_base,
                               size_t slot_bytes, int n_slots, int n_copy_threads) {
  RK_REQUIRE(out && pinned_base && device_base, RK_ERR_INVALID, "rk_stage_create: null argument");
  RK_REQUIRE(n_slots >= 2 && n_slots <= 64, RK_ERR_INVALID,
             "rk_stage_create: n_slots=%d outside [2,64]", n_slots);
  RK_REQUIRE(slot_bytes > 0 && slot_bytes % 256 == 0, RK_ERR_INVALID,
             "rk_stage_create: slot_bytes must be a positive multiple of 256");
  rk_stage* st = new rk_stage();
  st->pinned = (char*)pinned_base;
  st->dev = (char*)device_base;
  st->slot_bytes = slot_bytes;
  st->n_slots = n_slots;
  st->n_threads = n_copy_threads < 1 ? 1 : (n_copy_threads > 16 ? 16 : n_copy_threads);
  cudaError_t e = cudaGetDevice(&st->device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st->copy_stream, cudaStreamNonBlocking);
  st->copied.resize(n_slots);
  st->released.resize(n_slots);
  st->copied_valid.assign(n_slots, 0);
  st->released_valid.assign(n_slots, 0);
  for (int i = 0; i < n_slots && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&st->copied[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&st->released[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    rk_set_error("rk_stage_create: %s", cudaGetErrorString(e));
    delete st;
    return (int)e;
  }
  memset(st->ticket_slot, 0, sizeof(st->ticket_slot));
  st->worker = std::thread([st] { st->run(); });
  *out = st;
  return 0;
}

extern "C" int rk_stage_submit(rk_stage* st, int slot, const void* const* src,
                               const size_t* offsets, const size_t* sizes, int n_segments,
                               uint64_t* ticket) {
  RK_REQUIRE(st && ticket, RK_ERR_INVALID, "rk_stage_submit: null argument");
  RK_REQUIRE(slot >= 0 && slot < st->n_slots, RK_ERR_INVALID, "rk_stage_submit: bad slot %d", slot);
  RK_REQUIRE(n_segments >= 0 && (n_segments == 0 || (src && offsets && sizes)), RK_ERR_INVALID,
             "rk_stage_submit: null segment arrays");
  Job job;
  job.slot = slot;
  job.extent = 0;
  job.segs.reserve(n_segments);
  for (int i = 0; i < n_segments; ++i) {
    RK_REQUIRE(sizes[i] == 0 || src[i], RK_ERR_INVALID, "rk_stage_submit: segment %d is null", i);
    RK_REQUIRE(offsets[i] + sizes[i] <= st->slot_bytes, RK_ERR_TOO_LARGE,
               "rk_stage_submit: segment %d ends at %zu, slot holds %zu bytes", i,
               offsets[i] + sizes[i], st->slot_bytes);
    job.segs.push_back(Segment{(const char*)src[i], offsets[i], sizes[i]});
    if (offsets[i] + sizes[i] > job.extent) job.extent = offsets[i] + sizes[i];
  }
  {
    std::lock_guard<std::mutex> lk(st->mu);
    RK_REQUIRE(!st->err, st->err, "stager worker failed earlier: %s", st->err_msg.c_str());
    job.ticket = st->next_ticket++;
    st->ticket_slot[job.ticket % 64] = (uint64_t)slot;
    *ticket = job.ticket;
    st->q.push_back(std::move(job));
  }
  st->cv_job.notify_one();
  return 0;
}

extern "C" int rk_stage_wait(rk_stage* st, uint64_t ticket, void* consumer_stream) {
  RK_REQUIRE(st, RK_ERR_INVALID, "rk_stage_wait: null stager");
  int slot;
  {
    std::unique_lock<std::mutex> lk(st->mu);
    RK_REQUIRE(ticket >= 1 && ticket < st->next_ticket, RK_ERR_INVALID,
               "rk_stage_wait: unknown ticket %llu", (unsigned long long)ticket);
    st->cv_done.wait(lk, [&] { return st->issued >= ticket || st->err; });
    RK_REQUIRE(!st->err, st->err, "stager worker failed: %s", st->err_msg.c_str());
    slot = (int)st->ticket_slot[ticket % 64];
  }
  RK_CUDA(cudaStreamWaitEvent((cudaStream_t)consumer_stream, st->copied[slot], 0));
  return 0;
}

extern "C" int rk_stage_release(rk_stage* st, int slot, void* consumer_stream) {
  RK_REQUIRE(st, RK_ERR_INVALID, "rk_stage_release: null stager");
  RK_REQUIRE(slot >= 0 && slot < st->n_slots, RK_ERR_INVALID, "rk_stage_release: bad slot %d", slot);
  RK_CUDA(cudaEventRecord(st->released[slot], (cudaStream_t)consumer_stream));
  std::lock_guard<std::mutex> lk(st->mu);
  st->released_valid[slot] = 1;
  return 0;
}

extern "C" int rk_stage_destroy(rk_stage* st) {
  if (!st) return 0;
  {
    std::lock_guard<std::mutex> lk(st->mu);
    st->stop = true;
  }
  st->cv_job.notify_all();
  if (st->worker.joinable()) st->worker.join();
  cudaStreamSynchronize(st->copy_stream);
  for (auto e : st->copied) cudaEventDestroy(e);
  for (auto e : st->released) cudaEventDestroy(e);
  cudaStreamDestroy(st->copy_stream);
  delete st;
  return 0;
}
