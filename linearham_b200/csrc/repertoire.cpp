// Repertoire driver: many independent clonal families scored on one GPU (BASELINE config 5).
//
// The reference scores one family per process: SCons instantiates one `linearham --run-linearham`
// command per selected cluster (SConstruct:344-352, 428-447; src/linearham.cpp:377-386), every
// process building its own PhyloHMM and libpll partitions. Here a repertoire owns a few *slots* on
// one device - each a device handle (lh_family_t) with its streams, planner threads, pinned staging
// and device buffers - and one host thread per slot pulls families off a shared queue, binds the
// family to its handle (lh_family_rebind: one blob copy, no allocation once the buffers have grown
// to the largest family) and runs lh_score_batch. With two or more slots the planning, copies and
// small kernels of one family overlap the kernels of another, which is what keeps the GPU busy on
// families of 20-500 leaves whose single launches do not fill it.
//
// Only the C ABI of include/linearham_b200.h is used; there is no likelihood arithmetic here.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/linearham_b200.h"

struct lh_repertoire {
  std::vector<int> slot_device;  // the device of every slot (slots of the devices interleaved)
  int host_threads = 0;  // planner threads per slot
  std::vector<lh_family_t*> slots;
  std::vector<std::pair<std::string, int64_t>> options;  // lh_set_option on every slot handle
};

namespace {
thread_local std::string g_rep_err;
int rep_fail(const std::string& m) {
  g_rep_err = m;
  return 1;
}
// pruning cost model of sharding.py / SURVEY.md 8(e): trees x inner nodes x xMSA columns x categories
double item_cost(const lh_repertoire_item_t& it) {
  if (!it.desc || !it.trees) return 0.0;
  return (double)it.trees->n_trees * std::max(1, it.desc->n_tips - 2) * it.desc->n_sites * it.trees->n_rates;
}
}  // namespace

extern "C" {

const char* lh_repertoire_last_error(void) { return g_rep_err.c_str(); }

int lh_repertoire_create_multi(lh_repertoire_t** out, const int* devices, int n_devices, int n_slots, int host_threads) {
  if (!out || !devices) return rep_fail("null argument");
  *out = nullptr;
  int ndev = 0;
  if (lh_device_count(&ndev) || ndev == 0)
    return rep_fail(std::string("no CUDA device available; linearham_b200 has no CPU fallback (") + lh_last_error() + ")");
  if (n_devices < 1 || n_devices > 64) return rep_fail("need 1..64 devices");
  for (int i = 0; i < n_devices; ++i)
    if (devices[i] < 0 || devices[i] >= ndev) return rep_fail("device " + std::to_string(devices[i]) + " out of range");
  if (n_slots < 1 || n_slots > 16) return rep_fail("n_slots must be 1..16");
  lh_repertoire* r = new lh_repertoire;
  if (host_threads <= 0) {
    const int hw = std::max(1u, std::thread::hardware_concurrency());
    host_threads = std::max(1, hw / (n_slots * n_devices));
  }
  r->host_threads = host_threads;
  r->slots.assign((size_t)n_slots * n_devices, nullptr);  // handles are created by the first family a slot scores
  for (int si = 0; si < n_slots * n_devices; ++si) r->slot_device.push_back(devices[si % n_devices]);
  *out = r;
  return 0;
}

int lh_repertoire_create(lh_repertoire_t** out, int device, int n_slots, int host_threads) {
  return lh_repertoire_create_multi(out, &device, 1, n_slots, host_threads);
}

// lh_set_option on every slot handle, present and future ("host_threads" also sets the per-slot share).
int lh_repertoire_set_option(lh_repertoire_t* r, const char* key, int64_t value) {
  if (!r || !key) return rep_fail("null argument");
  for (lh_family_t* f : r->slots)
    if (f && lh_set_option(f, key, value)) return rep_fail(lh_last_error());
  if (std::string(key) == "host_threads") r->host_threads = (int)value;
  else r->options.emplace_back(key, value);
  return 0;
}

void lh_repertoire_destroy(lh_repertoire_t* r) {
  if (!r) return;
  for (lh_family_t* f : r->slots) lh_family_destroy(f);
  delete r;
}

int lh_repertoire_score(lh_repertoire_t* r, lh_repertoire_item_t* items, int32_t n_items,
                        lh_repertoire_stats_t* stats) {
  if (!r || (n_items > 0 && !items)) return rep_fail("null argument");
  if (n_items < 0) return rep_fail("negative family count");
  const auto t0 = std::chrono::steady_clock::now();
  // longest first, so that the slots finish together
  std::vector<int> order(n_items);
  std::iota(order.begin(), order.end(), 0);
  std::vector<double> cost(n_items);
  for (int i = 0; i < n_items; ++i) {
    cost[i] = item_cost(items[i]);
    items[i].status = -1;
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });

  std::atomic<int> next(0), failed(0);
  std::mutex err_mu;
  std::string first_err;
  struct SlotSum {
    int64_t launches = 0, h2d = 0, d2h = 0, trees = 0;
    double ms_bind = 0, ms_score = 0;
  };
  std::vector<SlotSum> sums(r->slots.size());
  auto work = [&](int si) {
    SlotSum& sm = sums[si];
    for (;;) {
      const int k = next.fetch_add(1);
      if (k >= n_items) break;
      lh_repertoire_item_t& it = items[order[k]];
      auto bad = [&](const std::string& m) {
        it.status = 1;
        failed.fetch_add(1);
        std::lock_guard<std::mutex> g(err_mu);
        if (first_err.empty()) first_err = "family " + std::to_string(order[k]) + ": " + m;
      };
      if (!it.desc || !it.trees || (!it.loglik && it.trees->n_trees > 0)) {
        bad("null descriptor, tree batch or output");
        continue;
      }
      const auto ta = std::chrono::steady_clock::now();
      lh_family_t*& f = r->slots[si];
      int rc;
      if (!f) {
        rc = lh_family_create(&f, it.desc, r->slot_device[si]);
        if (!rc) rc = lh_set_option(f, "host_threads", r->host_threads);
        // The slots share the host. With two or more host threads per slot the planner threads keep
        // ahead of the GPU and k_plan launches of a few hundred trees only add latency (C5, 16 cores, 6
        // slots: 247 ms per pass with the host planner, 260 ms with k_plan); with one thread per slot
        // (8 ranks on one host) the host is the bottleneck and the device plans (C5 on 4 cores: 297 ms
        // vs 279 ms). "device_plan" can still be set through lh_repertoire_set_option.
        if (!rc) rc = lh_set_option(f, "device_plan", r->host_threads <= 1 ? 1 : 0);
        for (const auto& kv : r->options)
          if (!rc) rc = lh_set_option(f, kv.first.c_str(), kv.second);
      } else {
        rc = lh_family_rebind(f, it.desc);
      }
      const auto tb = std::chrono::steady_clock::now();
      sm.ms_bind += std::chrono::duration<double, std::milli>(tb - ta).count();
      if (rc) {
        bad(lh_last_error());
        continue;
      }
      if (lh_score_batch(f, it.trees, it.loglik, nullptr)) {
        bad(lh_last_error());
        continue;
      }
      sm.ms_score += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb).count();
      lh_timings_t tm;
      if (!lh_get_timings(f, &tm)) {
        sm.launches += tm.kernel_launches;
        sm.h2d += tm.h2d_bytes;
        sm.d2h += tm.d2h_bytes;
      }
      sm.trees += it.trees->n_trees;
      it.status = 0;
    }
  };
  const int n_workers = (int)std::min<size_t>(r->slots.size(), (size_t)std::max(n_items, 1));
  std::vector<std::thread> threads;
  for (int si = 1; si < n_workers; ++si) threads.emplace_back(work, si);
  work(0);
  for (auto& t : threads) t.join();

  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    stats->n_families = n_items;
    stats->n_failed = failed.load();
    stats->n_slots = (int32_t)r->slots.size();
    stats->host_threads_per_slot = r->host_threads;
    for (const SlotSum& s : sums) {
      stats->n_trees += s.trees;
      stats->kernel_launches += s.launches;
      stats->h2d_bytes += s.h2d;
      stats->d2h_bytes += s.d2h;
      stats->ms_bind += (float)s.ms_bind;
      stats->ms_score += (float)s.ms_score;
    }
    stats->ms_wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  if (failed.load()) return rep_fail(first_err);
  return 0;
}

}  // extern "C"
