// The parameter server's apply daemon ("PS process ... parks, optionally running a
// persistent apply-server"). In the reference the PS task's CPU executes the ApplyAdam ops that
// replica_device_setter placed on /job:ps whenever a worker's gradient arrives
// (run_mnist_distributed.py:107,116; the PS itself just sits in server.join(), :96-98). Here the
// task that owns a shard arena runs ONE native host thread per shard that
//   1. polls the shard's mailbox flags (a 512-byte D2H copy on its own stream — copy engine only,
//      no SM is occupied while nothing is pending),
//   2. when some worker has published a gradient, launches dtf_mailbox_drain (snapshot + apply
//      from local HBM) on that stream, concurrently with whatever the colocated worker computes.
// No Python and no GIL on this path; kernels never wait on other kernels (the drain only touches
// gradients whose sequence number was already published), so nothing can deadlock on the device.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <thread>

#include "../kernels/common.cuh"
#include "../../include/dtf_b200.h"

using namespace dtf;

struct dtf_mbox_daemon {
  int device = 0;
  int opt = 0;
  dtf_apply_item items[DTF_MAX_APPLY_ITEMS];
  dtf_drain_mirror_item mirror[DTF_MAX_APPLY_ITEMS];
  bool has_mirror = false;
  int shard = 0;
  std::atomic<int> backstop_us{0};
  std::atomic<int> max_ctas{0};   // grid cap of the drains (0 = one CTA per chunk)
  int sync_R = 0;                 // > 0: synchronous rounds (replicas_to_aggregate), see dtf_mailbox_drain_sync
  unsigned int* host_epoch = nullptr;   // pinned
  int n_items = 0;
  long long slot_stride = 0;
  int depth = 1;
  int n_workers = 1;
  dtf_apply_hyper hyper;
  dtf_ps_ctrl* ctrl_powers = nullptr;
  dtf_ps_ctrl* ctrl_step = nullptr;
  dtf_mbox_flags* flags = nullptr;
  int poll_us = 20;

  cudaStream_t stream = nullptr;
  dtf_mbox_flags* host_flags = nullptr;  // pinned
  unsigned int* scratch = nullptr;       // device: todo[16] + lock-held word | block counter at +32
  std::thread thread;
  std::atomic<int> stop{0};
  std::atomic<int> paused{0};
  std::atomic<int> idle{0};
  std::atomic<int> failed{0};
  std::atomic<unsigned long long> polls{0}, drains{0}, applied{0};
  unsigned int last_consumed[DTF_MAX_WORKERS] = {0};
  // backstop mode: when worker w's oldest pending gradient was first seen, and at which consumed[w]
  unsigned int seen_consumed[DTF_MAX_WORKERS] = {0};
  std::chrono::steady_clock::time_point seen_at[DTF_MAX_WORKERS];
  bool seen[DTF_MAX_WORKERS] = {false};
  std::mutex err_mu;
  char err[512] = {0};

  void fail(const char* fmt, ...) {
    std::lock_guard<std::mutex> g(err_mu);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, sizeof err, fmt, ap);
    va_end(ap);
    failed.store(1);
  }
};

namespace {

int pending_count(const dtf_mbox_flags* f, int n_workers) {
  int n = 0;
  for (int w = 0; w < n_workers; ++w)
    if (static_cast<int>(f->produced[w] - f->consumed[w]) > 0) ++n;
  return n;
}

void daemon_main(dtf_mbox_daemon* d) {
  if (cudaSetDevice(d->device) != cudaSuccess) {
    d->fail("apply daemon: cudaSetDevice(%d) failed", d->device);
    return;
  }
  while (!d->stop.load(std::memory_order_acquire)) {
    if (d->paused.load(std::memory_order_acquire)) {
      d->idle.store(1, std::memory_order_release);
      std::this_thread::sleep_for(std::chrono::microseconds(50));
      continue;
    }
    d->idle.store(0, std::memory_order_release);
    cudaError_t e = cudaMemcpyAsync(d->host_flags, d->flags, sizeof(dtf_mbox_flags),
                                    cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess && d->sync_R > 0)
      e = cudaMemcpyAsync(d->host_epoch, &d->ctrl_powers->epoch, sizeof(unsigned int),
                          cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    if (e != cudaSuccess) {
      d->fail("apply daemon: polling the mailbox flags failed: %s", cudaGetErrorString(e));
      return;
    }
    d->polls.fetch_add(1, std::memory_order_relaxed);
    // exact accounting: gradients retired since the previous poll (a drain may retire more than the
    // poll that triggered it saw; a worker that re-attaches resets its counters to zero)
    for (int w = 0; w < d->n_workers; ++w) {
      const unsigned int c = d->host_flags->consumed[w];
      const int delta = d->polls.load(std::memory_order_relaxed) > 1 ? static_cast<int>(c - d->last_consumed[w]) : 0;
      if (delta > 0) d->applied.fetch_add(static_cast<unsigned long long>(delta), std::memory_order_relaxed);
      d->last_consumed[w] = c;
    }
    int n = pending_count(d->host_flags, d->n_workers);
    const int backstop = d->backstop_us.load(std::memory_order_relaxed);
    if (backstop > 0) {
      // the shard's drains normally run inside a colocated worker's step graph: step in only for
      // gradients that have been waiting for `backstop` microseconds (that worker has stopped)
      const auto now = std::chrono::steady_clock::now();
      bool overdue = false;
      for (int w = 0; w < d->n_workers; ++w) {
        const unsigned int c = d->host_flags->consumed[w];
        if (static_cast<int>(d->host_flags->produced[w] - c) <= 0) {
          d->seen[w] = false;
          continue;
        }
        if (!d->seen[w] || d->seen_consumed[w] != c) {
          d->seen[w] = true;
          d->seen_consumed[w] = c;
          d->seen_at[w] = now;
        } else if (std::chrono::duration_cast<std::chrono::microseconds>(now - d->seen_at[w]).count() >= backstop) {
          overdue = true;
        }
      }
      if (!overdue) {
        std::this_thread::sleep_for(std::chrono::microseconds(std::max(d->poll_us, backstop / 4)));
        continue;
      }
    }
    if (d->sync_R > 0) {
      // a round is actionable when enough FRESH gradients are pending, or a stale one waits to be retired
      const unsigned int epoch = *d->host_epoch;
      int fresh = 0, stale = 0;
      for (int w = 0; w < d->n_workers; ++w) {
        const unsigned int c = d->host_flags->consumed[w];
        if (static_cast<int>(d->host_flags->produced[w] - c) <= 0) continue;
        const unsigned int tag = (c & 1u) ? d->host_flags->tag1[w] : d->host_flags->tag0[w];
        if (tag == epoch) ++fresh;
        else if (static_cast<int>(tag - epoch) < 0) ++stale;
      }
      if (fresh < d->sync_R && stale == 0) {
        if (d->poll_us > 0) std::this_thread::sleep_for(std::chrono::microseconds(d->poll_us));
        continue;
      }
      if (dtf_mailbox_drain_sync(d->opt, d->items, d->n_items, d->slot_stride, d->n_workers, d->sync_R,
                                 &d->hyper, d->ctrl_powers, d->ctrl_step, d->flags, d->scratch,
                                 d->scratch + 32, d->has_mirror ? d->mirror : nullptr, d->shard,
                                 d->max_ctas.load(std::memory_order_relaxed), d->stream) != 0) {
        d->fail("apply daemon: %s", dtf_last_error());
        return;
      }
      e = cudaStreamSynchronize(d->stream);
      if (e != cudaSuccess) {
        d->fail("apply daemon: sync drain kernel failed: %s", cudaGetErrorString(e));
        return;
      }
      d->drains.fetch_add(1, std::memory_order_relaxed);
      continue;
    }
    if (n == 0) {
      if (d->poll_us > 0) std::this_thread::sleep_for(std::chrono::microseconds(d->poll_us));
      continue;
    }
    if (dtf_mailbox_drain_ex(d->opt, d->items, d->n_items, d->slot_stride, d->depth, d->n_workers,
                             &d->hyper, d->ctrl_powers, d->ctrl_step, d->flags, d->scratch,
                             d->scratch + 32, nullptr, d->has_mirror ? d->mirror : nullptr, d->shard,
                             d->max_ctas.load(std::memory_order_relaxed), d->stream) != 0) {
      d->fail("apply daemon: %s", dtf_last_error());
      return;
    }
    e = cudaStreamSynchronize(d->stream);
    if (e != cudaSuccess) {
      d->fail("apply daemon: drain kernel failed: %s", cudaGetErrorString(e));
      return;
    }
    d->drains.fetch_add(1, std::memory_order_relaxed);
  }
}

}  // namespace

extern "C" int dtf_mbox_daemon_start(dtf_mbox_daemon** out, int device, int opt,
                                     const dtf_apply_item* items, int n_items, long long slot_stride,
                                     int depth, int n_workers, const dtf_apply_hyper* hyper,
                                     dtf_ps_ctrl* ctrl_powers, dtf_ps_ctrl* ctrl_step,
                                     dtf_mbox_flags* flags, int poll_us,
                                     const dtf_drain_mirror_item* mirror_items, int shard_index,
                                     int backstop_us, int sync_replicas_to_aggregate) {
  DTF_REQUIRE(out != nullptr && hyper != nullptr && flags != nullptr, "mbox_daemon_start: null argument");
  DTF_REQUIRE(n_items >= 0 && n_items <= DTF_MAX_APPLY_ITEMS, "mbox_daemon_start: %d items (max %d)",
              n_items, DTF_MAX_APPLY_ITEMS);
  DTF_REQUIRE(depth >= 1 && n_workers >= 1 && n_workers <= DTF_MAX_WORKERS,
              "mbox_daemon_start: depth %d, %d workers", depth, n_workers);
  int prev = -1;
  DTF_CHECK_CUDA(cudaGetDevice(&prev));
  DTF_CHECK_CUDA(cudaSetDevice(device));
  dtf_mbox_daemon* d = new dtf_mbox_daemon();
  d->device = device;
  d->opt = opt;
  for (int i = 0; i < n_items; ++i) d->items[i] = items[i];
  d->has_mirror = mirror_items != nullptr;
  for (int i = 0; i < n_items && mirror_items != nullptr; ++i) d->mirror[i] = mirror_items[i];
  d->shard = shard_index;
  d->backstop_us.store(backstop_us > 0 ? backstop_us : 0);
  d->sync_R = sync_replicas_to_aggregate > 0 ? sync_replicas_to_aggregate : 0;
  DTF_REQUIRE(d->sync_R == 0 || (ctrl_powers != nullptr && depth <= 2),
              "mbox_daemon_start: synchronous mode needs the shard's control block and depth <= 2");
  d->n_items = n_items;
  d->slot_stride = slot_stride;
  d->depth = depth;
  d->n_workers = n_workers;
  d->hyper = *hyper;
  d->ctrl_powers = ctrl_powers;
  d->ctrl_step = ctrl_step;
  d->flags = flags;
  d->poll_us = poll_us;
  int lo = 0, hi = 0;
  cudaError_t e = cudaDeviceGetStreamPriorityRange(&lo, &hi);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&d->stream, cudaStreamNonBlocking, lo);
  if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&d->host_flags), sizeof(dtf_mbox_flags));
  if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&d->host_epoch), 64);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&d->scratch), (DTF_MAX_WORKERS + 48) * 4);
  if (e == cudaSuccess) e = cudaMemset(d->scratch, 0, (DTF_MAX_WORKERS + 48) * 4);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("mbox_daemon_start: %s", cudaGetErrorString(e));
    if (d->scratch) cudaFree(d->scratch);
    if (d->host_flags) cudaFreeHost(d->host_flags);
    if (d->host_epoch) cudaFreeHost(d->host_epoch);
    if (d->stream) cudaStreamDestroy(d->stream);
    delete d;
    cudaSetDevice(prev);
    return -1;
  }
  d->thread = std::thread(daemon_main, d);
  cudaSetDevice(prev);
  *out = d;
  return 0;
}

extern "C" int dtf_mbox_daemon_set_max_ctas(dtf_mbox_daemon* d, int max_ctas) {
  DTF_REQUIRE(d != nullptr, "mbox_daemon_set_max_ctas: null daemon");
  d->max_ctas.store(max_ctas > 0 ? max_ctas : 0, std::memory_order_relaxed);
  return 0;
}

extern "C" int dtf_mbox_daemon_set_backstop(dtf_mbox_daemon* d, int backstop_us) {
  DTF_REQUIRE(d != nullptr, "mbox_daemon_set_backstop: null daemon");
  d->backstop_us.store(backstop_us > 0 ? backstop_us : 0, std::memory_order_relaxed);
  return 0;
}

/* pause = 1 returns once the daemon is parked between drains (checkpoint / restore want a quiet
 * shard); pause = 0 resumes. */
extern "C" int dtf_mbox_daemon_pause(dtf_mbox_daemon* d, int pause) {
  DTF_REQUIRE(d != nullptr, "mbox_daemon_pause: null daemon");
  d->paused.store(pause ? 1 : 0, std::memory_order_release);
  if (pause) {
    for (int i = 0; i < 200000 && !d->idle.load(std::memory_order_acquire) && !d->failed.load(); ++i)
      std::this_thread::sleep_for(std::chrono::microseconds(10));
    DTF_REQUIRE(d->idle.load() || d->failed.load(), "mbox_daemon_pause: the daemon did not park");
  }
  return 0;
}

extern "C" int dtf_mbox_daemon_stats(dtf_mbox_daemon* d, unsigned long long* polls,
                                     unsigned long long* drains, unsigned long long* applied) {
  DTF_REQUIRE(d != nullptr, "mbox_daemon_stats: null daemon");
  if (polls) *polls = d->polls.load();
  if (drains) *drains = d->drains.load();
  if (applied) *applied = d->applied.load();
  if (d->failed.load()) {
    std::lock_guard<std::mutex> g(d->err_mu);
    set_error("%s", d->err);
    return -1;
  }
  return 0;
}

extern "C" int dtf_mbox_daemon_stop(dtf_mbox_daemon* d) {
  if (d == nullptr) return 0;
  d->stop.store(1, std::memory_order_release);
  if (d->thread.joinable()) d->thread.join();
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(d->device);
  int rc = 0;
  if (d->failed.load()) {
    set_error("%s", d->err);
    rc = -1;
  }
  cudaStreamDestroy(d->stream);
  cudaFreeHost(d->host_flags);
  cudaFreeHost(d->host_epoch);
  cudaFree(d->scratch);
  if (prev >= 0) cudaSetDevice(prev);
  delete d;
  return rc;
}
