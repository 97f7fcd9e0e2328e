// C ABI of include/linearham_b200.h: device memory management, host-side planning,
// stream orchestration. No arithmetic of the likelihood path lives here except the two
// libm constants log(2^256) / log(2^-256) the reference also takes from the host libm.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/linearham_b200.h"
#include "kernels.hpp"
#include "plan.hpp"

namespace {

thread_local std::string g_err;

int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

#define LH_CUDA(expr)                                                                    \
  do {                                                                                   \
    cudaError_t e_ = (expr);                                                             \
    if (e_ != cudaSuccess)                                                               \
      return fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// growable device / pinned-host buffers
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 4;
    LH_CUDA(cudaMalloc(&p, want));
    cap = want;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 4;
    LH_CUDA(cudaMallocHost(&p, want));
    cap = want;
    return 0;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

template <class T>
int upload(const T* src, size_t n, const T** dst, std::vector<void*>& owned) {
  void* d = nullptr;
  size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
  LH_CUDA(cudaMalloc(&d, bytes));
  owned.push_back(d);
  if (n) LH_CUDA(cudaMemcpy(d, src, n * sizeof(T), cudaMemcpyHostToDevice));
  *dst = reinterpret_cast<const T*>(d);
  return 0;
}

// dense column-major [rows x cols] -> CSC on the device (rows ascending inside a column,
// i.e. the order of the reference's dense dot product with the zero terms dropped)
int upload_csc(const double* dense, int rows, int cols, lhb::CscDev* out, std::vector<void*>& owned) {
  std::vector<int32_t> colptr(cols + 1, 0), row;
  std::vector<double> val;
  for (int j = 0; j < cols; ++j) {
    for (int i = 0; i < rows; ++i) {
      double v = dense[(size_t)j * rows + i];
      if (v != 0.0) {
        row.push_back(i);
        val.push_back(v);
      }
    }
    colptr[j + 1] = (int32_t)row.size();
  }
  out->rows = rows;
  out->cols = cols;
  if (upload(colptr.data(), colptr.size(), &out->colptr, owned)) return 1;
  if (upload(row.data(), row.size(), &out->row, owned)) return 1;
  if (upload(val.data(), val.size(), &out->val, owned)) return 1;
  return 0;
}

}  // namespace

struct lh_family {
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  lhb::FamilyDev fd{};
  std::vector<void*> owned;  // family-constant device allocations
  // options
  int opt_sites_per_thread = 0, opt_smem_slots = -1, opt_timing = 1, opt_threads = 0;
  size_t opt_chunk_bytes = 48u << 20;
  // batch state
  int B = 0, R = 0, slots = 0;
  DevBuf d_plan, d_pblen, d_er, d_pi, d_rates, d_eig, d_prec, d_xe, d_loglik, d_scratch, d_scratch_cnt;
  DevBuf d_vgerm, d_vd, d_dgerm, d_dj, d_jgerm, d_gs, d_vds, d_djs;
  PinBuf h_plan, h_pblen, h_misc, h_out;
  cudaEvent_t ev[8] = {};
  cudaEvent_t user_ev[8] = {};
  std::vector<cudaEvent_t> chunk_ev;  // 3 per chunk: pmatrix start, prune start, prune end
  int n_chunks = 0;
  bool own_stream = true;
  std::vector<uint8_t> h_tip_codes;  // host copy, re-tiled per sites_per_thread
  DevBuf d_codes_t;
  int codes_ns = 0;
  lh_timings_t tm{};
  bool want_tables = false;
};

extern "C" {

const char* lh_last_error(void) { return g_err.c_str(); }
const char* lh_version(void) { return "linearham_b200 0.1 (sm_100a)"; }

int lh_device_count(int* n) {
  *n = 0;
  LH_CUDA(cudaGetDeviceCount(n));
  return 0;
}

void lh_family_destroy(lh_family_t* f) {
  if (!f) return;
  cudaSetDevice(f->device);
  if (f->stream) cudaStreamSynchronize(f->stream);
  for (void* p : f->owned) cudaFree(p);
  DevBuf* bufs[] = {&f->d_plan, &f->d_pblen, &f->d_er, &f->d_pi, &f->d_rates, &f->d_eig, &f->d_prec,
                    &f->d_xe, &f->d_loglik, &f->d_scratch, &f->d_scratch_cnt, &f->d_vgerm, &f->d_vd,
                    &f->d_dgerm, &f->d_dj, &f->d_jgerm, &f->d_gs, &f->d_vds, &f->d_djs, &f->d_codes_t};
  for (DevBuf* b : bufs) b->release();
  f->h_plan.release();
  f->h_pblen.release();
  f->h_misc.release();
  f->h_out.release();
  for (auto& e : f->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : f->user_ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : f->chunk_ev) cudaEventDestroy(e);
  if (f->stream && f->own_stream) cudaStreamDestroy(f->stream);
  delete f;
}

int lh_family_create(lh_family_t** out, const lh_family_desc_t* d, int device) {
  *out = nullptr;
  if (!d) return fail("null descriptor");
  if (d->n_tips < 3) return fail("need at least 3 tips (got %d)", d->n_tips);
  if (d->n_sites < 1) return fail("empty xMSA");
  if (d->n_v < 1 || d->n_j < 1 || (d->has_d && d->n_d < 1)) return fail("empty germline region");
  if (d->vd_sites < 1 || d->vd_states < 1 || (d->has_d && (d->dj_sites < 1 || d->dj_states < 1)))
    return fail("empty junction region");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail("no CUDA device available (%s); linearham_b200 has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail("device %d out of range (0..%d)", device, ndev - 1);
  LH_CUDA(cudaSetDevice(device));
  lh_family* f = new lh_family;
  f->device = device;
  cudaDeviceGetAttribute(&f->num_sms, cudaDevAttrMultiProcessorCount, device);
  auto bail = [&](int) { lh_family_destroy(f); return 1; };
  if (cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail("cudaStreamCreate failed"));
  for (auto& ev : f->ev) cudaEventCreate(&ev);
  for (auto& ev : f->user_ev) cudaEventCreate(&ev);

  const int T = d->n_tips, S = d->n_sites;
  for (size_t i = 0; i < (size_t)T * S; ++i)
    if (d->tip_codes[i] > 4) return bail(fail("tip_codes[%zu] = %d is not in 0..4", i, d->tip_codes[i]));
  lhb::FamilyDev& fd = f->fd;
  fd.T = T;
  fd.S = S;
  fd.has_d = d->has_d ? 1 : 0;
  fd.n_v = d->n_v;
  fd.n_d = fd.has_d ? d->n_d : 0;
  fd.n_j = d->n_j;
  fd.vd_sites = d->vd_sites;
  fd.vd_states = d->vd_states;
  fd.dj_sites = fd.has_d ? d->dj_sites : 0;
  fd.dj_states = fd.has_d ? d->dj_states : 0;
  fd.g_max = std::max(fd.n_v, std::max(fd.n_d, fd.n_j));
  fd.s_max = std::max(std::max(fd.vd_states, fd.dj_states), fd.g_max);
  auto& own = f->owned;
  int rc = 0;
  rc |= upload(d->tip_codes, (size_t)T * S, &fd.tip_codes, own);
  f->h_tip_codes.assign(d->tip_codes, d->tip_codes + (size_t)T * S);
  auto region = [&](int n, const int32_t* off, const int32_t* inds, const int32_t** doff,
                    const int32_t** dinds, const char* name) {
    if (n == 0) return upload(off, 0, doff, own) | upload(inds, 0, dinds, own);
    for (int g = 0; g < n; ++g)
      if (off[g + 1] < off[g]) return fail("%s offsets not monotone", name);
    for (int j = off[0]; j < off[n]; ++j)
      if (inds[j] < 0 || inds[j] >= S) return fail("%s xMSA index %d out of range", name, inds[j]);
    if (off[0] != 0) return fail("%s offsets must start at 0", name);
    return upload(off, n + 1, doff, own) | upload(inds, off[n], dinds, own);
  };
  rc |= region(fd.n_v, d->vpad_off, d->vpad_inds, &fd.vpad_off, &fd.vpad_inds, "vpad");
  rc |= region(fd.n_v, d->vgerm_off, d->vgerm_inds, &fd.vgerm_off, &fd.vgerm_inds, "vgerm");
  rc |= region(fd.n_d, d->dgerm_off, d->dgerm_inds, &fd.dgerm_off, &fd.dgerm_inds, "dgerm");
  rc |= region(fd.n_j, d->jgerm_off, d->jgerm_inds, &fd.jgerm_off, &fd.jgerm_inds, "jgerm");
  rc |= region(fd.n_j, d->jpad_off, d->jpad_inds, &fd.jpad_off, &fd.jpad_inds, "jpad");
  if (rc) return bail(1);
  auto junction = [&](const int32_t* cm, int L, int Sj, const int32_t** dst) {
    std::vector<int32_t> rm((size_t)L * Sj);
    for (int i = 0; i < L; ++i)
      for (int j = 0; j < Sj; ++j) {
        int32_t v = cm[(size_t)j * L + i];
        if (v < -1 || v >= S) return fail("junction xMSA index %d out of range", v);
        rm[(size_t)i * Sj + j] = v;
      }
    return upload(rm.data(), rm.size(), dst, own);
  };
  rc |= junction(d->vd_inds, fd.vd_sites, fd.vd_states, &fd.vd_inds_rm);
  rc |= junction(d->dj_inds, fd.dj_sites, fd.dj_states, &fd.dj_inds_rm);
  rc |= upload(d->v_gene_prob, fd.n_v, &fd.v_gene_prob, own);
  rc |= upload(d->v_trans_prod, fd.n_v, &fd.v_trans_prod, own);
  rc |= upload(d->vpad_trans, fd.n_v, &fd.vpad_trans, own);
  rc |= upload(d->jpad_trans, fd.n_j, &fd.jpad_trans, own);
  rc |= upload_csc(d->t_v_vd, fd.n_v, fd.vd_states, &fd.t_v_vd, own);
  rc |= upload_csc(d->t_vd, fd.vd_states, fd.vd_states, &fd.t_vd, own);
  rc |= upload_csc(d->t_vd_d, fd.vd_states, fd.has_d ? fd.n_d : fd.n_j, &fd.t_vd_d, own);
  if (fd.has_d) {
    rc |= upload_csc(d->t_d_dj, fd.n_d, fd.dj_states, &fd.t_d_dj, own);
    rc |= upload_csc(d->t_dj, fd.dj_states, fd.dj_states, &fd.t_dj, own);
    rc |= upload_csc(d->t_dj_j, fd.dj_states, fd.n_j, &fd.t_dj_j, own);
  }
  if (rc) return bail(1);
  *out = f;
  return 0;
}

int lh_set_option(lh_family_t* f, const char* key, int64_t value) {
  std::string k(key ? key : "");
  if (k == "sites_per_thread") {
    if (value < 0 || value > 4) return fail("sites_per_thread must be 0 (auto) or 1..4");
    f->opt_sites_per_thread = (int)value;
  } else if (k == "smem_slots") {
    f->opt_smem_slots = (int)value;
  } else if (k == "timing") {
    f->opt_timing = value != 0;
  } else if (k == "host_threads") {
    f->opt_threads = (int)value;
  } else if (k == "chunk_bytes") {
    f->opt_chunk_bytes = (size_t)value;
  } else if (k == "stream") {
    // run on a caller-owned CUDA stream (e.g. torch.cuda.current_stream().cuda_stream); 0 restores
    // a private stream
    cudaSetDevice(f->device);
    if (f->stream) cudaStreamSynchronize(f->stream);
    if (f->stream && f->own_stream) cudaStreamDestroy(f->stream);
    if (value == 0) {
      f->own_stream = true;
      if (cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking) != cudaSuccess) return fail("cudaStreamCreate failed");
    } else {
      f->own_stream = false;
      f->stream = reinterpret_cast<cudaStream_t>(static_cast<intptr_t>(value));
    }
  } else {
    return fail("unknown option '%s'", k.c_str());
  }
  return 0;
}

}  // extern "C"

namespace {

// xMSA columns per thread: fewest padded columns over the whole xMSA; ties go to the smaller
// register footprint.
int pick_sites_per_thread(int S, int threads) {
  if (S <= 4 * threads) return std::max(1, (S + threads - 1) / threads);
  // 4 columns per thread amortise the per-step work best; fall back to 3 only when that pads
  // the xMSA by a quarter less
  const int pad4 = ((S + 4 * threads - 1) / (4 * threads)) * 4 * threads;
  const int pad3 = ((S + 3 * threads - 1) / (3 * threads)) * 3 * threads;
  return (pad3 * 4 < pad4 * 3) ? 3 : 4;
}

// Plans all trees into the pinned staging buffers (multi-threaded); returns max slots.
int plan_batch(lh_family* f, const lh_tree_batch_t* tb, lhb::Step* steps, double* pblen, int* max_slots) {
  const int T = f->fd.T, B = tb->n_trees, n_ops = T - 2, NB = 2 * T - 3;
  int nthreads = f->opt_threads > 0 ? f->opt_threads : (int)std::thread::hardware_concurrency();
  nthreads = std::max(1, std::min(nthreads, std::min(32, (B + 255) / 256)));
  std::atomic<int> bad(-1);
  std::vector<int> tmax(nthreads, 0);
  auto work = [&](int tid) {
    lhb::PlanScratch sc;
    int mx = 0;
    for (int t = tid; t < B; t += nthreads) {
      int s = lhb::plan_tree(T, tb->children + (size_t)t * n_ops * 2, tb->blen + (size_t)t * n_ops * 2,
                             tb->root_child[t], tb->root_blen[t], steps + (size_t)t * (T - 1),
                             pblen + (size_t)t * NB, sc);
      if (s < 0) {
        bad.store(t);
        return;
      }
      mx = std::max(mx, s);
    }
    tmax[tid] = mx;
  };
  if (nthreads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; ++i) th.emplace_back(work, i);
    for (auto& t : th) t.join();
  }
  if (bad.load() >= 0) return fail("tree %d is not a valid post-order operation list", bad.load());
  *max_slots = *std::max_element(tmax.begin(), tmax.end());
  return 0;
}

}  // namespace

extern "C" {

int lh_batch_upload(lh_family_t* f, const lh_tree_batch_t* tb) {
  if (!f || !tb) return fail("null argument");
  const int B = tb->n_trees, R = tb->n_rates, T = f->fd.T, S = f->fd.S, NB = 2 * T - 3;
  if (B < 0) return fail("negative batch size");
  if (R < 1 || R > LH_MAX_RATES) return fail("n_rates %d outside 1..%d", R, LH_MAX_RATES);
  LH_CUDA(cudaSetDevice(f->device));
  f->B = B;
  f->R = R;
  f->tm = lh_timings_t{};
  f->tm.n_trees = B;
  if (B == 0) return 0;
  for (int t = 0; t < B; ++t) {
    for (int i = 0; i < 4; ++i)
      if (!(tb->pi[4 * t + i] > 0.0)) return fail("tree %d: pi[%d] must be positive", t, i);
    for (int i = 0; i < NB - 1; ++i)
      if (!(tb->blen[(size_t)t * (NB - 1) + i] >= 0.0)) return fail("tree %d: negative branch length", t);
    if (!(tb->root_blen[t] >= 0.0)) return fail("tree %d: negative branch length", t);
  }
  const size_t plan_bytes = (size_t)B * (T - 1) * sizeof(lhb::Step);
  const size_t pblen_bytes = (size_t)B * NB * sizeof(double);
  const size_t misc_doubles = (size_t)B * (6 + 4 + R);
  if (f->h_plan.reserve(plan_bytes) || f->h_pblen.reserve(pblen_bytes) ||
      f->h_misc.reserve(misc_doubles * sizeof(double)))
    return 1;
  const auto plan_t0 = std::chrono::steady_clock::now();
  if (plan_batch(f, tb, f->h_plan.as<lhb::Step>(), f->h_pblen.as<double>(), &f->slots)) return 1;
  f->tm.ms_plan = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - plan_t0).count();
  double* misc = f->h_misc.as<double>();
  std::memcpy(misc, tb->er, sizeof(double) * 6 * B);
  std::memcpy(misc + (size_t)6 * B, tb->pi, sizeof(double) * 4 * B);
  std::memcpy(misc + (size_t)10 * B, tb->rates, sizeof(double) * R * B);
  if (f->d_plan.reserve(plan_bytes) || f->d_pblen.reserve(pblen_bytes) ||
      f->d_er.reserve(misc_doubles * sizeof(double)) || f->d_eig.reserve((size_t)B * lhb::kEigStride * 8) ||
      f->d_xe.reserve((size_t)B * S * 8) || f->d_loglik.reserve((size_t)B * 8))
    return 1;
  if (f->opt_timing) cudaEventRecord(f->ev[0], f->stream);
  LH_CUDA(cudaMemcpyAsync(f->d_plan.p, f->h_plan.p, plan_bytes, cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaMemcpyAsync(f->d_pblen.p, f->h_pblen.p, pblen_bytes, cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaMemcpyAsync(f->d_er.p, f->h_misc.p, misc_doubles * 8, cudaMemcpyHostToDevice, f->stream));
  if (f->opt_timing) cudaEventRecord(f->ev[1], f->stream);
  f->tm.h2d_bytes = (int64_t)(plan_bytes + pblen_bytes + misc_doubles * 8);
  return 0;
}

int lh_batch_run(lh_family_t* f) {
  if (!f) return fail("null family");
  if (f->B == 0) return 0;
  LH_CUDA(cudaSetDevice(f->device));
  const int B = f->B, R = f->R, T = f->fd.T, S = f->fd.S, NB = 2 * T - 3;
  lhb::BatchDev bd{};
  bd.B = B;
  bd.R = R;
  bd.plan = f->d_plan.as<int4>();
  bd.pblen = f->d_pblen.as<double>();
  bd.er = f->d_er.as<double>();
  bd.pi = bd.er + (size_t)6 * B;
  bd.rates = bd.er + (size_t)10 * B;
  bd.eig = f->d_eig.as<double>();
  bd.xe = f->d_xe.as<double>();
  bd.loglik = f->d_loglik.as<double>();

  // pruning configuration
  lhb::PruneCfg pc{};
  pc.sites_per_thread = f->opt_sites_per_thread ? f->opt_sites_per_thread : pick_sites_per_thread(S, 32);
  const int tile_sites = 32 * pc.sites_per_thread;
  pc.n_tiles = (S + tile_sites - 1) / tile_sites;
  pc.total_slots = f->slots;
  const size_t slot_bytes = (size_t)tile_sites * (4 * sizeof(double) + sizeof(int));
  int occ0 = lhb::prune_max_ctas_per_sm(pc.sites_per_thread, 0);
  if (occ0 < 1) return fail("pruning kernel cannot be resident (sites/thread=%d)", pc.sites_per_thread);
  int smem_slots = f->opt_smem_slots;
  if (smem_slots < 0) {
    size_t per_cta = (size_t)(227 * 1024) / occ0 - 1024 - lhb::prune_smem_bytes(pc.sites_per_thread, 0);
    smem_slots = (int)(per_cta / slot_bytes);
  }
  smem_slots = std::max(0, std::min(smem_slots, pc.total_slots));
  while (smem_slots > 0 && lhb::prune_smem_bytes(pc.sites_per_thread, smem_slots) > (size_t)227 * 1024) --smem_slots;
  pc.smem_slots = smem_slots;
  int occ = lhb::prune_max_ctas_per_sm(pc.sites_per_thread, pc.smem_slots);
  if (occ < 1) return fail("pruning kernel cannot be resident with %d shared-memory slots", pc.smem_slots);
  pc.log_threshold = std::log(1.0 / std::pow(2.0, 256));
  if (f->codes_ns != pc.sites_per_thread) {
    // re-tile the tip codes for this NS (once per family and NS)
    const int NS = pc.sites_per_thread;
    std::vector<uint32_t> ct((size_t)T * pc.n_tiles * 32, 0u);
    for (int r = 0; r < T; ++r)
      for (int tile = 0; tile < pc.n_tiles; ++tile)
        for (int lane = 0; lane < 32; ++lane) {
          uint32_t word = 0;
          for (int u = 0; u < NS; ++u) {
            const int site = tile * 32 * NS + u * 32 + lane;
            if (site < S) word |= (uint32_t)(f->h_tip_codes[(size_t)r * S + site] * 32u) << (8 * u);
          }
          ct[((size_t)r * pc.n_tiles + tile) * 32 + lane] = word;
        }
    if (f->d_codes_t.reserve(ct.size() * 4)) return 1;
    LH_CUDA(cudaMemcpyAsync(f->d_codes_t.p, ct.data(), ct.size() * 4, cudaMemcpyHostToDevice, f->stream));
    LH_CUDA(cudaStreamSynchronize(f->stream));
    f->codes_ns = NS;
  }
  pc.codes_t = f->d_codes_t.as<uint32_t>();

  // chunk the batch so one chunk's branch records stay L2-resident
  const size_t rec_bytes_tree = (size_t)NB * 20 * R * sizeof(double);
  int chunk = (int)std::max<size_t>(1, f->opt_chunk_bytes / rec_bytes_tree);
  const int resident = occ * f->num_sms;
  // at least 16 waves of work items per launch when the batch allows it, so the partial last
  // wave costs a few percent. The records then exceed L2 for large trees, which is harmless: they
  // are written once and read once per site tile by CTAs that run at the same time (the tiles
  // of a tree are adjacent work items), ~2.6 MB of DRAM traffic per 1001-tip tree.
  chunk = std::max(chunk, (16 * resident + pc.n_tiles - 1) / pc.n_tiles);
  chunk = std::min(chunk, B);
  if (f->d_prec.reserve((size_t)chunk * rec_bytes_tree)) return 1;
  bd.prec = f->d_prec.as<double>();
  pc.grid = std::min(resident, chunk * pc.n_tiles);
  const int gslots = pc.total_slots - pc.smem_slots;
  if (gslots > 0) {
    if (f->d_scratch.reserve((size_t)pc.grid * gslots * 4 * tile_sites * sizeof(double)) ||
        f->d_scratch_cnt.reserve((size_t)pc.grid * gslots * tile_sites * sizeof(int)))
      return 1;
  }
  pc.scratch = f->d_scratch.as<double>();
  pc.scratch_cnt = f->d_scratch_cnt.as<int32_t>();

  lhb::HmmOut ho{};
  ho.log_factor = std::log(std::pow(2.0, 256));
  if (f->want_tables) {
    const lhb::FamilyDev& fd = f->fd;
    if (f->d_vgerm.reserve((size_t)B * fd.n_v * 8) || f->d_vd.reserve((size_t)B * fd.vd_sites * fd.vd_states * 8) ||
        f->d_dgerm.reserve((size_t)B * std::max(fd.n_d, 1) * 8) ||
        f->d_dj.reserve((size_t)B * std::max(fd.dj_sites * fd.dj_states, 1) * 8) ||
        f->d_jgerm.reserve((size_t)B * fd.n_j * 8) || f->d_gs.reserve((size_t)B * 3 * 4) ||
        f->d_vds.reserve((size_t)B * fd.vd_sites * 4) || f->d_djs.reserve((size_t)B * std::max(fd.dj_sites, 1) * 4))
      return 1;
    ho.vgerm_forward = f->d_vgerm.as<double>();
    ho.vd_forward = f->d_vd.as<double>();
    ho.dgerm_forward = fd.has_d ? f->d_dgerm.as<double>() : nullptr;
    ho.dj_forward = fd.has_d ? f->d_dj.as<double>() : nullptr;
    ho.jgerm_forward = f->d_jgerm.as<double>();
    ho.germ_scalers = f->d_gs.as<int32_t>();
    ho.vd_scalers = f->d_vds.as<int32_t>();
    ho.dj_scalers = fd.has_d ? f->d_djs.as<int32_t>() : nullptr;
  }

  cudaStream_t s = f->stream;
  int launches = 0;
  const int n_chunks = (B + chunk - 1) / chunk;
  f->n_chunks = n_chunks;
  if (f->opt_timing)
    while ((int)f->chunk_ev.size() < 3 * n_chunks) {
      cudaEvent_t e;
      LH_CUDA(cudaEventCreate(&e));
      f->chunk_ev.push_back(e);
    }
  if (f->opt_timing) cudaEventRecord(f->ev[2], s);
  lhb::launch_eigen(bd, s);
  ++launches;
  for (int t0 = 0, ci = 0; t0 < B; t0 += chunk, ++ci) {
    const int n = std::min(chunk, B - t0);
    if (f->opt_timing) cudaEventRecord(f->chunk_ev[3 * ci], s);
    lhb::launch_pmatrix(f->fd, bd, t0, n, s);
    if (f->opt_timing) cudaEventRecord(f->chunk_ev[3 * ci + 1], s);
    lhb::PruneCfg c = pc;
    c.grid = std::min(pc.grid, n * pc.n_tiles);
    lhb::launch_prune(f->fd, bd, c, t0, n, s);
    if (f->opt_timing) cudaEventRecord(f->chunk_ev[3 * ci + 2], s);
    launches += 2;
  }
  if (f->opt_timing) cudaEventRecord(f->ev[3], s);
  lhb::launch_hmm(f->fd, bd, ho, s);
  ++launches;
  if (f->opt_timing) cudaEventRecord(f->ev[4], s);
  LH_CUDA(cudaGetLastError());
  f->tm.kernel_launches = launches;
  f->tm.prune_launches = n_chunks;
  f->tm.prune_slots = pc.total_slots;
  f->tm.prune_tile_sites = tile_sites;
  f->tm.prune_sites_per_thread = pc.sites_per_thread;
  f->tm.prune_smem_slots = pc.smem_slots;
  f->tm.prune_grid = pc.grid;
  f->tm.prune_ctas_per_sm = occ;
  return 0;
}

// Per-kernel device times of the last lh_batch_run (syncs the stream).
static int collect_run_timings(lh_family* f) {
  if (!f->opt_timing || f->B == 0) return 0;
  LH_CUDA(cudaEventSynchronize(f->ev[4]));
  float pm = 0, pr = 0, x = 0;
  for (int ci = 0; ci < f->n_chunks; ++ci) {
    cudaEventElapsedTime(&x, f->chunk_ev[3 * ci], f->chunk_ev[3 * ci + 1]);
    pm += x;
    cudaEventElapsedTime(&x, f->chunk_ev[3 * ci + 1], f->chunk_ev[3 * ci + 2]);
    pr += x;
  }
  f->tm.ms_pmatrix = pm;
  f->tm.ms_prune = pr;
  cudaEventElapsedTime(&f->tm.ms_hmm, f->ev[3], f->ev[4]);
  cudaEventElapsedTime(&f->tm.ms_kernels, f->ev[2], f->ev[4]);
  return 0;
}

int lh_sync(lh_family_t* f) {
  if (!f) return fail("null family");
  LH_CUDA(cudaSetDevice(f->device));
  LH_CUDA(cudaStreamSynchronize(f->stream));
  return 0;
}

int lh_loglik_dev(lh_family_t* f, const double** p, int32_t* n) {
  if (!f) return fail("null family");
  *p = f->d_loglik.as<double>();
  *n = f->B;
  return 0;
}

int lh_batch_fetch(lh_family_t* f, double* loglik, lh_outputs_t* o) {
  if (!f) return fail("null family");
  const int B = f->B;
  if (B == 0) return 0;
  LH_CUDA(cudaSetDevice(f->device));
  const lhb::FamilyDev& fd = f->fd;
  cudaStream_t s = f->stream;
  if (f->opt_timing) cudaEventRecord(f->ev[5], s);
  int64_t bytes = 0;
  auto get = [&](void* dst, const void* src, size_t n) {
    if (!dst || !n) return 0;
    bytes += (int64_t)n;
    LH_CUDA(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s));
    return 0;
  };
  if (f->h_out.reserve((size_t)B * 8)) return 1;
  if (get(f->h_out.p, f->d_loglik.p, (size_t)B * 8)) return 1;
  if (o) {
    if (get(o->xmsa_emission, f->d_xe.p, (size_t)B * fd.S * 8)) return 1;
    const bool tables = o->vgerm_forward || o->vd_forward || o->dgerm_forward || o->dj_forward ||
                        o->jgerm_forward || o->germ_scalers || o->vd_scalers || o->dj_scalers;
    if (tables) {
      if (!f->want_tables) return fail("forward tables were not produced by the last run (internal)");
      if (get(o->vgerm_forward, f->d_vgerm.p, (size_t)B * fd.n_v * 8) ||
          get(o->vd_forward, f->d_vd.p, (size_t)B * fd.vd_sites * fd.vd_states * 8) ||
          get(o->jgerm_forward, f->d_jgerm.p, (size_t)B * fd.n_j * 8) ||
          get(o->germ_scalers, f->d_gs.p, (size_t)B * 3 * 4) ||
          get(o->vd_scalers, f->d_vds.p, (size_t)B * fd.vd_sites * 4))
        return 1;
      if (fd.has_d && (get(o->dgerm_forward, f->d_dgerm.p, (size_t)B * fd.n_d * 8) ||
                       get(o->dj_forward, f->d_dj.p, (size_t)B * fd.dj_sites * fd.dj_states * 8) ||
                       get(o->dj_scalers, f->d_djs.p, (size_t)B * fd.dj_sites * 4)))
        return 1;
    }
  }
  if (f->opt_timing) cudaEventRecord(f->ev[6], s);
  LH_CUDA(cudaStreamSynchronize(s));
  LH_CUDA(cudaGetLastError());
  std::memcpy(loglik, f->h_out.p, (size_t)B * 8);
  f->tm.d2h_bytes = bytes;
  if (f->opt_timing) {
    collect_run_timings(f);
    cudaEventElapsedTime(&f->tm.ms_h2d, f->ev[0], f->ev[1]);
    cudaEventElapsedTime(&f->tm.ms_d2h, f->ev[5], f->ev[6]);
    cudaEventElapsedTime(&f->tm.ms_total, f->ev[0], f->ev[6]);
  }
  return 0;
}

int lh_score_batch(lh_family_t* f, const lh_tree_batch_t* tb, double* loglik, lh_outputs_t* o) {
  if (!f || !tb || !loglik) return fail("null argument");
  f->want_tables = o && (o->vgerm_forward || o->vd_forward || o->dgerm_forward || o->dj_forward ||
                         o->jgerm_forward || o->germ_scalers || o->vd_scalers || o->dj_scalers);
  if (lh_batch_upload(f, tb)) return 1;
  if (lh_batch_run(f)) return 1;
  return lh_batch_fetch(f, loglik, o);
}

int lh_measure_fp64_peak(int device, double* tflops) {
  if (!tflops) return fail("null argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return fail("no such CUDA device %d", device);
  LH_CUDA(cudaSetDevice(device));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const int blocks = sms * 8;
  double* scratch = nullptr;
  LH_CUDA(cudaMalloc(&scratch, (size_t)blocks * 256 * sizeof(double)));
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  double best = 0.0;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(a, 0);
    const double flops = lhb::fp64_peak_flops_per_launch(scratch, blocks, 4096, 0);
    cudaEventRecord(b, 0);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (rep && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(scratch);
  LH_CUDA(cudaGetLastError());
  *tflops = best;
  return 0;
}

int lh_get_timings(lh_family_t* f, lh_timings_t* out) {
  if (!f || !out) return fail("null argument");
  LH_CUDA(cudaSetDevice(f->device));
  if (collect_run_timings(f)) return 1;
  *out = f->tm;
  return 0;
}

int lh_timer_mark(lh_family_t* f, int32_t slot) {
  if (!f || slot < 0 || slot >= 8) return fail("timer slot out of range");
  LH_CUDA(cudaSetDevice(f->device));
  LH_CUDA(cudaEventRecord(f->user_ev[slot], f->stream));
  return 0;
}

int lh_timer_elapsed_ms(lh_family_t* f, int32_t a, int32_t b, float* ms) {
  if (!f || !ms || a < 0 || a >= 8 || b < 0 || b >= 8) return fail("timer slot out of range");
  LH_CUDA(cudaSetDevice(f->device));
  LH_CUDA(cudaEventSynchronize(f->user_ev[b]));
  LH_CUDA(cudaEventElapsedTime(ms, f->user_ev[a], f->user_ev[b]));
  return 0;
}

}  // extern "C"
