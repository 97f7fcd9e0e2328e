// Multi-GPU handle: one clonal family on several devices of one box, inside one process (SURVEY 8(b)/8(e):
// `const int* devices, int n_devices`). The path shards with no exchange - the trees of a batch are cut into
// contiguous, equal shards, one per device; every member is an ordinary lh_family_t on its device with its own
// streams, planner threads and staging; a call runs the members on one host thread each and the results land in
// the caller's single host arrays (plain per-GPU copies: there is no reduction to make). Which device scored a
// tree does not change its doubles: the members run the same kernels on the same inputs
// (tests/test_gpu_group.py: bit-identical to one GPU).
//
// Only the C ABI of include/linearham_b200.h is used; there is no likelihood arithmetic here.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/linearham_b200.h"

struct lh_group {
  std::vector<lh_family_t*> members;
  std::vector<int> devices;
  // per-tree strides of the optional read-outs
  int T = 0, S = 0, n_v = 0, n_d = 0, n_j = 0, vd_n = 0, dj_n = 0, vd_sites = 0, dj_sites = 0;
  std::vector<int32_t> last_lo, last_n;  // the shards of the last call
};

namespace {
thread_local std::string g_grp_err;
int grp_fail(const std::string& m) {
  g_grp_err = m;
  return 1;
}

// "tree 5: ..." of a member is tree lo + 5 of the caller's batch
std::string rebase_tree(const std::string& msg, int lo) {
  int idx = 0, used = 0;
  if (lo && std::sscanf(msg.c_str(), "tree %d%n", &idx, &used) == 1) return "tree " + std::to_string(idx + lo) + msg.substr(used);
  return msg;
}

template <class T>
T* shift(T* p, size_t n) { return p ? p + n : nullptr; }

lh_outputs_t shard_outputs(const lh_group* g, const lh_outputs_t* o, size_t lo, int R) {
  lh_outputs_t s{};
  if (!o) return s;
  s.xmsa_emission = shift(o->xmsa_emission, lo * g->S);
  s.vgerm_forward = shift(o->vgerm_forward, lo * g->n_v);
  s.vd_forward = shift(o->vd_forward, lo * g->vd_n);
  s.dgerm_forward = shift(o->dgerm_forward, lo * std::max(g->n_d, 0));
  s.dj_forward = shift(o->dj_forward, lo * g->dj_n);
  s.jgerm_forward = shift(o->jgerm_forward, lo * g->n_j);
  s.germ_scalers = shift(o->germ_scalers, lo * 3);
  s.vd_scalers = shift(o->vd_scalers, lo * g->vd_sites);
  s.dj_scalers = shift(o->dj_scalers, lo * g->dj_sites);
  s.site_rate_loglik = shift(o->site_rate_loglik, lo * (size_t)R * g->S);
  return s;
}

// runs fn(member index) on one host thread per member (the caller's thread takes member 0); first error wins
template <class F>
int run_members(lh_group* g, int n_active, F fn) {
  std::vector<std::string> err(n_active);
  std::vector<int> rc(n_active, 0);
  auto work = [&](int m) {
    rc[m] = fn(m);
    if (rc[m]) err[m] = lh_last_error();   // (thread-local in the member's thread)
  };
  std::vector<std::thread> th;
  for (int m = 1; m < n_active; ++m) th.emplace_back(work, m);
  work(0);
  for (auto& t : th) t.join();
  for (int m = 0; m < n_active; ++m)
    if (rc[m]) return grp_fail(rebase_tree(err[m], g->last_lo[m]));
  return 0;
}

void cut(lh_group* g, int B) {
  const int n = (int)g->members.size(), per = (B + n - 1) / n;
  g->last_lo.assign(n, 0);
  g->last_n.assign(n, 0);
  for (int m = 0; m < n; ++m) {
    const int lo = std::min(B, m * per);
    g->last_lo[m] = lo;
    g->last_n[m] = std::min(B, lo + per) - lo;
  }
}
}  // namespace

extern "C" {

const char* lh_group_last_error(void) { return g_grp_err.c_str(); }

// "0-3", "0,2,5", "1" -> device ordinals (the --devices flag and LINEARHAM_DEVICES)
int lh_parse_devices(const char* spec, int* out, int cap, int* n_out) {
  if (!spec || !out || !n_out) return grp_fail("null argument");
  int n = 0;
  const char* p = spec;
  while (*p) {
    char* end = nullptr;
    const long a = std::strtol(p, &end, 10);
    if (end == p || a < 0) return grp_fail(std::string("cannot read the device list '") + spec + "'");
    long b = a;
    p = end;
    if (*p == '-') {
      b = std::strtol(p + 1, &end, 10);
      if (end == p + 1 || b < a) return grp_fail(std::string("cannot read the device list '") + spec + "'");
      p = end;
    }
    for (long d = a; d <= b; ++d) {
      if (n >= cap) return grp_fail("too many devices");
      for (int i = 0; i < n; ++i)
        if (out[i] == d) return grp_fail("device " + std::to_string(d) + " is listed twice");
      out[n++] = (int)d;
    }
    if (*p == ',') ++p;
    else if (*p) return grp_fail(std::string("cannot read the device list '") + spec + "'");
  }
  if (n == 0) return grp_fail("empty device list");
  *n_out = n;
  return 0;
}

void lh_group_destroy(lh_group_t* g) {
  if (!g) return;
  for (lh_family_t* f : g->members) lh_family_destroy(f);
  delete g;
}

int lh_group_create(lh_group_t** out, const lh_family_desc_t* d, const int* devices, int n_devices) {
  if (!out || !d || !devices) return grp_fail("null argument");
  *out = nullptr;
  if (n_devices < 1 || n_devices > 64) return grp_fail("need 1..64 devices");
  for (int i = 0; i < n_devices; ++i)
    for (int j = 0; j < i; ++j)
      if (devices[i] == devices[j]) return grp_fail("device " + std::to_string(devices[i]) + " is listed twice");
  lh_group* g = new lh_group;
  g->devices.assign(devices, devices + n_devices);
  g->members.assign(n_devices, nullptr);
  g->T = d->n_tips;
  g->S = d->n_sites;
  g->n_v = d->n_v;
  g->n_d = d->has_d ? d->n_d : 0;
  g->n_j = d->n_j;
  g->vd_sites = d->vd_sites;
  g->vd_n = d->vd_sites * d->vd_states;
  g->dj_sites = d->has_d ? d->dj_sites : 0;
  g->dj_n = d->has_d ? d->dj_sites * d->dj_states : 0;
  g->last_lo.assign(n_devices, 0);
  g->last_n.assign(n_devices, 0);
  // the family constants go to every device (replicated), one thread per device
  const int hw = std::max(1u, std::thread::hardware_concurrency());
  const int share = std::max(1, hw / n_devices);
  const int rc = run_members(g, n_devices, [&](int m) {
    if (lh_family_create(&g->members[m], d, g->devices[m])) return 1;
    return lh_set_option(g->members[m], "host_threads", share);  // the members share the host's cores
  });
  if (rc) {
    lh_group_destroy(g);
    return 1;
  }
  *out = g;
  return 0;
}

int lh_group_size(const lh_group_t* g) { return g ? (int)g->members.size() : 0; }
lh_family_t* lh_group_member(lh_group_t* g, int m) {
  return (g && m >= 0 && m < (int)g->members.size()) ? g->members[m] : nullptr;
}

int lh_group_set_option(lh_group_t* g, const char* key, int64_t value) {
  if (!g || !key) return grp_fail("null argument");
  for (lh_family_t* f : g->members)
    if (lh_set_option(f, key, value)) return grp_fail(lh_last_error());
  return 0;
}

int lh_group_set_labels(lh_group_t* g, const char* const* labels) {
  if (!g || !labels) return grp_fail("null argument");
  return run_members(g, (int)g->members.size(), [&](int m) { return lh_family_set_labels(g->members[m], labels); });
}

int lh_group_score_batch(lh_group_t* g, const lh_tree_batch_t* tb, double* loglik, lh_outputs_t* o) {
  if (!g || !tb || (!loglik && tb->n_trees > 0)) return grp_fail("null argument");
  const int B = tb->n_trees, R = tb->n_rates;
  if (B < 0) return grp_fail("negative batch size");
  cut(g, B);
  const size_t per_tree = (size_t)(g->T - 2) * 2;
  return run_members(g, (int)g->members.size(), [&](int m) {
    const size_t lo = (size_t)g->last_lo[m];
    lh_tree_batch_t s = *tb;
    s.n_trees = g->last_n[m];
    if (s.n_trees == 0) return 0;
    s.children = tb->children + lo * per_tree;
    s.blen = tb->blen + lo * per_tree;
    s.root_child = tb->root_child + lo;
    s.root_blen = tb->root_blen + lo;
    s.er = tb->er + lo * 6;
    s.pi = tb->pi + lo * 4;
    s.rates = tb->rates + lo * R;
    lh_outputs_t so = shard_outputs(g, o, lo, R);
    return lh_score_batch(g->members[m], &s, loglik + lo, o ? &so : nullptr);
  });
}

int lh_group_score_newick(lh_group_t* g, const lh_newick_batch_t* nb, double* loglik, lh_outputs_t* o) {
  if (!g || !nb || (!loglik && nb->n_trees > 0)) return grp_fail("null argument");
  const int B = nb->n_trees, R = nb->n_rates;
  if (B < 0) return grp_fail("negative batch size");
  cut(g, B);
  return run_members(g, (int)g->members.size(), [&](int m) {
    const size_t lo = (size_t)g->last_lo[m];
    lh_newick_batch_t s = *nb;
    s.n_trees = g->last_n[m];
    if (s.n_trees == 0) return 0;
    s.newick = nb->newick + lo;
    s.length = shift(nb->length, lo);
    s.er = nb->er + lo * 6;
    s.pi = nb->pi + lo * 4;
    s.rates = shift(nb->rates, lo * R);
    s.alpha = shift(nb->alpha, lo);
    lh_outputs_t so = shard_outputs(g, o, lo, R);
    return lh_score_newick(g->members[m], &s, loglik + lo, o ? &so : nullptr);
  });
}

// Timings of the last call: member m's own record (m >= 0), or with m = -1 the sums over the members of the
// byte / launch / tree counts with the slowest member's times.
int lh_group_get_timings(lh_group_t* g, int m, lh_timings_t* out) {
  if (!g || !out) return grp_fail("null argument");
  if (m >= 0) {
    if (m >= (int)g->members.size()) return grp_fail("no such member");
    if (lh_get_timings(g->members[m], out)) return grp_fail(lh_last_error());
    return 0;
  }
  lh_timings_t sum{};
  for (size_t i = 0; i < g->members.size(); ++i) {
    if (g->last_n[i] == 0) continue;
    lh_timings_t t;
    if (lh_get_timings(g->members[i], &t)) return grp_fail(lh_last_error());
    if (t.ms_call >= sum.ms_call) {
      const int64_t h = sum.h2d_bytes, d = sum.d2h_bytes;
      const int l = sum.kernel_launches, pl = sum.prune_launches, n = sum.n_trees, pt = sum.pattern_trees, nh = sum.newick_host_trees, td = sum.text_direct;
      sum = t;
      sum.h2d_bytes = h; sum.d2h_bytes = d; sum.kernel_launches = l; sum.prune_launches = pl; sum.n_trees = n;
      sum.pattern_trees = pt; sum.newick_host_trees = nh; sum.text_direct = td;
    }
    sum.h2d_bytes += t.h2d_bytes;
    sum.d2h_bytes += t.d2h_bytes;
    sum.kernel_launches += t.kernel_launches;
    sum.prune_launches += t.prune_launches;
    sum.n_trees += t.n_trees;
    sum.pattern_trees += t.pattern_trees;
    sum.newick_host_trees += t.newick_host_trees;
    sum.text_direct += t.text_direct;
  }
  *out = sum;
  return 0;
}

}  // extern "C"
