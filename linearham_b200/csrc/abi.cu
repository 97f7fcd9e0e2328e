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
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <emmintrin.h>

#include "../../include/linearham_b200.h"
#include "../../include/linearham_b200_host.h"
#include "kernels.hpp"
#include "plan.hpp"
#include "pool.hpp"

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

// memcpy into a pinned staging buffer with non-temporal stores. Lines written with ordinary stores sit
// dirty in the caches of the host cores that wrote them, and the GPU's DMA read then runs at ~8 GB/s
// (measured: 6.7 MB in 0.8 ms); streamed lines go to memory and are read at the link rate.
inline void stream_copy(void* dst, const void* src, size_t bytes) {
  uint8_t* d = static_cast<uint8_t*>(dst);
  const uint8_t* s = static_cast<const uint8_t*>(src);
  if ((reinterpret_cast<uintptr_t>(d) & 15) != 0 || bytes < 64) {
    std::memcpy(d, s, bytes);
    return;
  }
  size_t i = 0;
  for (; i + 64 <= bytes; i += 64) {
    const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + i));
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + i + 16));
    const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + i + 32));
    const __m128i e = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + i + 48));
    _mm_stream_si128(reinterpret_cast<__m128i*>(d + i), a);
    _mm_stream_si128(reinterpret_cast<__m128i*>(d + i + 16), b);
    _mm_stream_si128(reinterpret_cast<__m128i*>(d + i + 32), c);
    _mm_stream_si128(reinterpret_cast<__m128i*>(d + i + 48), e);
  }
  if (i < bytes) std::memcpy(d + i, s + i, bytes - i);
  _mm_sfence();
}

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

// Family constants are laid out in ONE host blob and go to HBM in one copy into one device arena
// (the handle keeps both, so binding another family to the handle allocates nothing once the arena
// has grown to the largest family seen). add() records where the device address of each array goes.
struct ConstArena {
  std::vector<uint8_t> host;
  struct Fix { const void** where; size_t off; };
  std::vector<Fix> fixes;
  void clear() { host.clear(); fixes.clear(); }
  template <class T>
  int add(const T* src, size_t n, const T** dst) {
    const size_t off = (host.size() + 255) / 256 * 256;
    host.resize(off + std::max<size_t>(n, 1) * sizeof(T), 0);
    if (n) std::memcpy(host.data() + off, src, n * sizeof(T));
    fixes.push_back({reinterpret_cast<const void**>(dst), off});
    return 0;
  }
};

template <class T>
int upload(const T* src, size_t n, const T** dst, ConstArena& arena) {
  return arena.add(src, n, dst);
}

// dense column-major [rows x cols] -> CSC on the device (rows ascending inside a column,
// i.e. the order of the reference's dense dot product with the zero terms dropped)
int upload_csc(const double* dense, int rows, int cols, lhb::CscDev* out, ConstArena& owned) {
  std::vector<int32_t> colptr(cols + 1, 0);
  std::vector<double> ent;  // (value, row as the low word of a double) pairs
  for (int j = 0; j < cols; ++j) {
    for (int i = 0; i < rows; ++i) {
      double v = dense[(size_t)j * rows + i];
      if (v != 0.0) {
        double packed = 0.0;
        const int64_t row64 = i;
        std::memcpy(&packed, &row64, sizeof(packed));
        ent.push_back(v);
        ent.push_back(packed);
      }
    }
    colptr[j + 1] = (int32_t)(ent.size() / 2);
  }
  out->rows = rows;
  out->cols = cols;
  if (upload(colptr.data(), colptr.size(), &out->colptr, owned)) return 1;
  if (upload(ent.data(), ent.size(), reinterpret_cast<const double**>(&out->ent), owned)) return 1;
  return 0;
}

}  // namespace

constexpr int kChunkEvents = 9;  // [7], [8]: k_plan of the chunk started / finished (plan stream)
constexpr size_t kTextSlotBytes = (size_t)32 << 20;  // a pinned ring slot for Newick text (the longest string it takes)
// ... filled only this far unless one string is longer: the copy of a segment overlaps the staging of the next, and
// with few host threads per GPU (8 ranks on one host) a chunk's text is otherwise staged, then copied, in turn
constexpr size_t kTextSegBytes = (size_t)6 << 20;

struct lh_family {
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  lhb::FamilyDev fd{};
  ConstArena owned;   // family constants: host blob + address fix-ups
  DevBuf d_const;     // ... and their device arena
  PinBuf h_const;
  // options
  int opt_sites_per_thread = 0, opt_smem_slots = -1, opt_smem_lo = -1, opt_timing = 1, opt_threads = 0;
  bool window_set = false;  // smem_lo of the batch has been chosen
  struct WindowCache {      // ... and what lh_score_newick chose for this family last time
    bool valid = false;
    int opt_lo = -1, lo[2] = {0, 0}, smem_slots[2] = {0, 0}, total_slots[2] = {0, 0}, slots = 0;
  } win_cache;
  size_t opt_chunk_bytes = 48u << 20;
  int opt_chunk_trees = 0, opt_pipe_waves = 4;
  // batch state
  int B = 0, R = 0, slots = 0;
  DevBuf d_plan, d_pblen, d_er, d_pi, d_rates, d_eig, d_prec, d_xe, d_loglik, d_scratch, d_scratch_cnt;
  DevBuf d_vgerm, d_vd, d_dgerm, d_dj, d_jgerm, d_gs, d_vds, d_djs;
  PinBuf h_misc, h_out;
  // staging ring: planned steps + branch lengths of one segment of trees per slot; the copy
  // stream drains a slot while the host plans into the next one
  static constexpr int kRing = 3;
  PinBuf h_ring[kRing];
  cudaEvent_t ring_ev[kRing] = {};
  bool ring_busy[kRing] = {};
  cudaEvent_t misc_ev = nullptr, list_ev = nullptr;  // h_misc / h_list have been copied
  bool misc_busy = false, list_busy = false;
  cudaStream_t copy_stream = nullptr;
  // The forward pass of a chunk runs here, beside the pruning of the next chunk (it is latency-bound:
  // a chunk's worth of trees leaves most of every SM idle).
  cudaStream_t hmm_stream = nullptr;
  cudaEvent_t hmm_done = nullptr;
  std::vector<cudaEvent_t> edge_ev;  // per chunk: k_edge finished
  int opt_hmm_overlap = 1;
  DevBuf d_counters;
  // device planning (k_plan): the caller's operation lists go to HBM as they are
  int opt_device_plan = -1;    // -1 auto, 0 host planner threads, 1 k_plan
  bool dev_plan = false;       // the current batch is planned on the device
  DevBuf d_rchildren, d_rblen, d_rroot_child, d_rroot_blen, d_flags;
  PinBuf h_flags;              // [0..3] init values, [4..7] read-back
  cudaStream_t plan_stream = nullptr;  // k_plan of a chunk runs here, beside the kernels of the chunk before
  bool plan_in_call = false;           // the chunk events [7], [8] of the current batch bracket k_plan launches
  bool flags_pending = false;  // k_plan ran for the batch and its flags have not been checked yet
  bool run_recorded = false;   // the kernels of the current batch have been queued (their chunk events exist)
  cudaEvent_t ev[8] = {};
  cudaEvent_t user_ev[8] = {};
  std::vector<cudaEvent_t> chunk_ev;  // kChunkEvents per chunk: start, pmatrix | (unused) | prune | edge ends (compute stream); hmm start, hmm end (its stream)
  std::vector<cudaEvent_t> staged_ev;  // per chunk: its last segment has landed in HBM
  int n_chunks = 0;
  bool own_stream = true;
  bool bound = false;  // a family is bound (false after a failed rebind)
  // pruning column sets: [0] all S xMSA columns, [1] the distinct patterns of rows 1..T-1
  struct ColumnSet {
    int n_cols = 0;
    std::vector<uint8_t> codes;  // [T][n_cols] host copy, re-tiled per sites_per_thread
    DevBuf d_codes_t;
    PinBuf h_codes_t;
    int codes_ns = 0;
  } cols[2];
  const int32_t* d_col_of_site = nullptr;  // [S] pattern index of each xMSA column
  DevBuf d_ebuf, d_ecnt, d_enaive;
  int opt_site_patterns = 1;
  // launch configuration of the current batch (configure_batch)
  lhb::BatchDev bd{};
  lhb::PruneCfg pc[2] = {};  // per column set
  lhb::EdgeCfg ec{};
  lhb::HmmOut ho{};
  int chunk = 0, chunk_pipe = 0, wave_pipe = 0, wave_unit = 0, seg = 0, occ[2] = {0, 0};
  int opt_open_unit = 1, opt_chunk_growth = 1;
  int opt_parse_reserve = 0;  // pruning CTAs per SM given up for the parser while text is coming in (measured: no gain, the parser of a chunk runs between the pruning launches either way)
  int opt_asr_scratch_mb = 0; // lh_asr_sample: CLV scratch per chunk of trees (0 = 4096 MB)
  void* asr_cache = nullptr;  // lh_asr_sample's device pool, kept between calls (lhb::asr_cache_free)
  int opt_keep_tables = 0;    // forward tables stay on the device after a run although nobody fetches them (lh_sample_paths)
  DevBuf d_words, d_sidx, d_sflag;   // lh_sample_paths: raw generator outputs in, drawn indices and flags out
  int opt_text_direct = 1;    // lh_score_newick: copy text the caller holds in pinned memory in place (0: always stage it)
  bool text_direct_call = false;  // this call's text is pinned and comes with its lengths
  int opt_newick_kernel = 0;  // 0 / 2 k_newick_cta (one CTA per tree), 1 k_newick (one thread per tree), 3 k_newick_warp (one warp per tree)
  double stage_ratio = 1.0;
  size_t rec_bytes_tree = 1;  // branch records of one tree, all categories
  // trees that can run on site patterns (naive tip on the root edge or under the virtual root)
  std::vector<uint8_t> on_patterns;
  int n_pattern_trees = 0;
  std::vector<int> starts;         // chunk boundaries of the current batch
  std::vector<int> chunk_n_pat;    // per chunk: trees on patterns (listed first in d_list)
  DevBuf d_list;                   // [B] tree indices, per chunk: pattern trees, then the others
  PinBuf h_list;
  lh_timings_t tm{};
  bool want_tables = false, want_rate_ll = false;
  int bad_plan_tree = -1;  // set by lh_batch_fetch when k_plan refused a tree
  DevBuf d_rate_ll;
  // Newick input (lh_score_newick): the xMSA row labels as a device hash table, a ring of text segments,
  // and what k_newick / k_split leave per tree and per chunk
  std::vector<std::string> labels;
  std::vector<const char*> label_ptrs;
  DevBuf d_label_slots, d_label_off, d_label_chars;
  lhb::LabelTableDev lt{};
  bool labels_set = false;
  bool newick_mode = false;           // the current batch came in as text
  DevBuf d_textbuf[2], d_items, d_toffs, d_tlens;   // the text of a chunk (two alternate), parser scratch, string offsets / lengths
  cudaEvent_t text_copied = nullptr, parsed_ev[2] = {};
  DevBuf d_on_pat, d_nw_status, d_chunk_npat;
  PinBuf h_nw_status, h_chunk_npat;
  std::vector<cudaEvent_t> seg_ev;    // pairs around the k_newick launches of the current call
  int n_seg_ev = 0;
  std::unique_ptr<lhb::WorkerPool> pool;  // planner threads, started on first use
  std::vector<lhb::PlanScratch> plan_scratch;
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
  cudaStreamSynchronize(f->stream);  // (a null handle is the legacy default stream)
  if (f->copy_stream) cudaStreamSynchronize(f->copy_stream);
  if (f->hmm_stream) cudaStreamSynchronize(f->hmm_stream);
  DevBuf* bufs[] = {&f->d_const, &f->d_plan, &f->d_pblen, &f->d_er, &f->d_pi, &f->d_rates, &f->d_eig, &f->d_prec,
                    &f->d_xe, &f->d_loglik, &f->d_scratch, &f->d_scratch_cnt, &f->d_vgerm, &f->d_vd,
                    &f->d_dgerm, &f->d_dj, &f->d_jgerm, &f->d_gs, &f->d_vds, &f->d_djs, &f->cols[0].d_codes_t,
                    &f->cols[1].d_codes_t, &f->d_ebuf, &f->d_ecnt, &f->d_enaive, &f->d_list, &f->d_counters,
                    &f->d_rate_ll, &f->d_rchildren, &f->d_rblen, &f->d_rroot_child, &f->d_rroot_blen, &f->d_flags,
                    &f->d_label_slots, &f->d_label_off, &f->d_label_chars, &f->d_textbuf[0], &f->d_textbuf[1], &f->d_items, &f->d_toffs, &f->d_tlens,
                    &f->d_on_pat, &f->d_nw_status, &f->d_chunk_npat, &f->d_words, &f->d_sidx, &f->d_sflag};
  for (DevBuf* b : bufs) b->release();
  lhb::asr_cache_free(f->asr_cache);
  f->asr_cache = nullptr;
  for (auto& r : f->h_ring) r.release();
  for (auto& c : f->cols) c.h_codes_t.release();
  f->h_misc.release();
  f->h_const.release();
  f->h_flags.release();
  f->h_list.release();
  f->h_out.release();
  f->h_nw_status.release();
  f->h_chunk_npat.release();
  if (f->text_copied) cudaEventDestroy(f->text_copied);
  for (auto& e : f->parsed_ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : f->seg_ev) cudaEventDestroy(e);
  for (auto& e : f->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : f->user_ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : f->ring_ev)
    if (e) cudaEventDestroy(e);
  if (f->misc_ev) cudaEventDestroy(f->misc_ev);
  if (f->list_ev) cudaEventDestroy(f->list_ev);
  for (auto& e : f->chunk_ev) cudaEventDestroy(e);
  for (auto& e : f->staged_ev) cudaEventDestroy(e);
  for (auto& e : f->edge_ev) cudaEventDestroy(e);
  if (f->hmm_done) cudaEventDestroy(f->hmm_done);
  if (f->hmm_stream) cudaStreamDestroy(f->hmm_stream);
  if (f->plan_stream) cudaStreamDestroy(f->plan_stream);
  if (f->copy_stream) cudaStreamDestroy(f->copy_stream);
  if (f->stream && f->own_stream) cudaStreamDestroy(f->stream);
  delete f;
}

static int check_desc(const lh_family_desc_t* d) {
  if (!d) return fail("null descriptor");
  if (d->n_tips < 3) return fail("need at least 3 tips (got %d)", d->n_tips);
  if (d->n_sites < 1) return fail("empty xMSA");
  if (d->n_v < 1 || d->n_j < 1 || (d->has_d && d->n_d < 1)) return fail("empty germline region");
  if (d->vd_sites < 1 || d->vd_states < 1 || (d->has_d && (d->dj_sites < 1 || d->dj_states < 1)))
    return fail("empty junction region");
  return 0;
}

// Binds the family described by `d` to the handle: builds every family constant into the handle's
// host blob and moves it to the handle's device arena in one copy. Batch buffers are kept.
static int bind_family(lh_family* f, const lh_family_desc_t* d) {
  LH_CUDA(cudaSetDevice(f->device));
  // nothing of an earlier family may still be running on the arena
  LH_CUDA(cudaStreamSynchronize(f->stream));
  LH_CUDA(cudaStreamSynchronize(f->copy_stream));
  LH_CUDA(cudaStreamSynchronize(f->hmm_stream));
  LH_CUDA(cudaStreamSynchronize(f->plan_stream));
  f->owned.clear();
  f->fd = lhb::FamilyDev{};
  f->B = 0;
  f->n_chunks = 0;
  f->window_set = false;
  f->labels_set = false;  // the labels belong to the family
  f->win_cache.valid = false;
  for (auto& c : f->cols) c.codes_ns = 0;
  auto bail = [&](int) { return 1; };
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
  {
    // site patterns: xMSA columns that agree on every row but the naive row (row 0) share all
    // CLVs away from the naive tip, so pruning visits each pattern once (first occurrence order)
    f->cols[0].n_cols = S;
    f->cols[0].codes.assign(d->tip_codes, d->tip_codes + (size_t)T * S);
    // (columns are compared through a transposed copy: contiguous keys, one pass over the matrix)
    std::vector<int32_t> col_of_site(S), rep;
    const int KT = T - 1;
    std::vector<uint8_t> tr((size_t)S * KT);
    for (int r0 = 1; r0 < T; r0 += 64)
      for (int c = 0; c < S; ++c)
        for (int r = r0; r < std::min(T, r0 + 64); ++r) tr[(size_t)c * KT + (r - 1)] = d->tip_codes[(size_t)r * S + c];
    std::unordered_map<uint64_t, std::vector<int32_t>> seen;  // column hash -> pattern ids with that hash
    seen.reserve((size_t)S * 2);
    for (int c = 0; c < S; ++c) {
      const uint8_t* key = &tr[(size_t)c * KT];
      uint64_t h = 1469598103934665603ull;
      int i = 0;
      for (; i + 8 <= KT; i += 8) {
        uint64_t w;
        std::memcpy(&w, key + i, 8);
        h = (h ^ w) * 1099511628211ull;
        h ^= h >> 29;
      }
      for (; i < KT; ++i) h = (h ^ key[i]) * 1099511628211ull;
      std::vector<int32_t>& bucket = seen[h];
      int32_t id = -1;
      for (int32_t cand : bucket)
        if (std::memcmp(&tr[(size_t)rep[cand] * KT], key, (size_t)KT) == 0) {
          id = cand;
          break;
        }
      if (id < 0) {
        id = (int32_t)rep.size();
        rep.push_back(c);
        bucket.push_back(id);
      }
      col_of_site[c] = id;
    }
    const int U = (int)rep.size();
    f->cols[1].n_cols = U;
    f->cols[1].codes.assign((size_t)T * U, 4);  // row 0 (the deferred naive tip) reads as N
    for (int r = 1; r < T; ++r)
      for (int u = 0; u < U; ++u) f->cols[1].codes[(size_t)r * U + u] = d->tip_codes[(size_t)r * S + rep[u]];
    rc |= upload(col_of_site.data(), (size_t)S, &f->d_col_of_site, own);
  }
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
  // per junction site, the states that can emit there (k_hmm skips the dot products of the others)
  auto emitting = [&](const int32_t* cm, int L, int Sj, const int32_t** doff, const int2** dpairs) {
    std::vector<int32_t> off(L + 1, 0), pairs;
    for (int i = 0; i < L; ++i) {
      for (int j = 0; j < Sj; ++j)
        if (cm[(size_t)j * L + i] >= 0) {
          pairs.push_back(j);
          pairs.push_back(cm[(size_t)j * L + i]);
        }
      off[i + 1] = (int32_t)(pairs.size() / 2);
    }
    if (pairs.empty()) pairs.assign(2, 0);
    return upload(off.data(), off.size(), doff, own) |
           upload(pairs.data(), pairs.size(), reinterpret_cast<const int32_t**>(dpairs), own);
  };
  rc |= emitting(d->vd_inds, fd.vd_sites, fd.vd_states, &fd.vd_valid_off, &fd.vd_valid);
  if (d->has_d) rc |= emitting(d->dj_inds, fd.dj_sites, fd.dj_states, &fd.dj_valid_off, &fd.dj_valid);
  {
    // a silent state's value is (finite dot product) * 0 = 0 as long as the dot product cannot overflow:
    // rows are checked on the device (entries <= 1e150), the matrices here (entries <= 1e100)
    bool small = true;
    auto scan = [&](const double* m, size_t n) {
      for (size_t i = 0; i < n; ++i) small &= (std::fabs(m[i]) <= 1e100);
    };
    scan(d->t_v_vd, (size_t)fd.n_v * fd.vd_states);
    scan(d->t_vd, (size_t)fd.vd_states * fd.vd_states);
    if (d->has_d) {
      scan(d->t_d_dj, (size_t)fd.n_d * fd.dj_states);
      scan(d->t_dj, (size_t)fd.dj_states * fd.dj_states);
    }
    fd.skip_silent = small ? 1 : 0;
  }
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
  // one copy of the blob; the recorded fields then get their device addresses
  const size_t bytes = own.host.size();
  if (f->d_const.reserve(bytes) || f->h_const.reserve(bytes)) return 1;
  stream_copy(f->h_const.p, own.host.data(), bytes);
  LH_CUDA(cudaMemcpyAsync(f->d_const.p, f->h_const.p, bytes, cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaStreamSynchronize(f->stream));
  for (const auto& fx : own.fixes) *fx.where = f->d_const.as<uint8_t>() + fx.off;
  f->bound = true;
  return 0;
}


int lh_family_create(lh_family_t** out, const lh_family_desc_t* d, int device) {
  if (!out) return fail("null argument");
  *out = nullptr;
  if (check_desc(d)) return 1;
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
  if (cudaStreamCreateWithFlags(&f->copy_stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail("cudaStreamCreate failed"));
  if (cudaStreamCreateWithFlags(&f->hmm_stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail("cudaStreamCreate failed"));
  if (cudaStreamCreateWithFlags(&f->plan_stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail("cudaStreamCreate failed"));
  cudaEventCreateWithFlags(&f->hmm_done, cudaEventDisableTiming);
  if (f->d_counters.reserve(64)) return bail(1);
  for (auto& ev : f->ring_ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&f->text_copied, cudaEventDisableTiming);
  for (auto& ev : f->parsed_ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&f->misc_ev, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&f->list_ev, cudaEventDisableTiming);
  for (auto& ev : f->ev) cudaEventCreate(&ev);
  for (auto& ev : f->user_ev) cudaEventCreate(&ev);
  if (bind_family(f, d)) return bail(1);
  *out = f;
  return 0;
}

// The handle takes on another family (same device); its streams, events, planner threads, pinned
// staging and device buffers are reused, so scoring many small families pays no allocation per family.
int lh_family_rebind(lh_family_t* f, const lh_family_desc_t* d) {
  if (!f) return fail("null family");
  if (check_desc(d)) return 1;
  f->bound = false;
  return bind_family(f, d);
}

// The xMSA row labels of the bound family -> the device hash table k_newick looks tips up in.
int lh_family_set_labels(lh_family_t* f, const char* const* labels) {
  if (!f || !labels) return fail("null argument");
  if (!f->bound) return fail("no family bound to the handle");
  LH_CUDA(cudaSetDevice(f->device));
  const int T = f->fd.T;
  if (T >= 65536) return fail("the device Newick parser takes families of up to 65535 tips (got %d)", T);
  f->labels_set = false;
  f->labels.assign(T, std::string());
  f->label_ptrs.assign(T, nullptr);
  size_t cap = 8;
  while (cap < (size_t)T * 2) cap <<= 1;
  std::vector<uint32_t> slots(cap * 4, 0xffffffffu), off(T, 0);
  std::vector<char> chars;
  for (int r = 0; r < T; ++r) {
    if (!labels[r]) return fail("null label for row %d", r);
    f->labels[r] = labels[r];
    const std::string& l = f->labels[r];
    if (l.empty() || l.size() >= 65536) return fail("label of row %d is empty or too long", r);
    uint32_t h = 2166136261u;  // FNV-1a, as host/newick.cpp
    uint64_t first8 = 0;
    for (size_t i = 0; i < l.size(); ++i) {
      h = (h ^ (unsigned char)l[i]) * 16777619u;
      if (i < 8) first8 |= (uint64_t)(unsigned char)l[i] << (8 * i);
    }
    off[r] = (uint32_t)chars.size();
    chars.insert(chars.end(), l.begin(), l.end());
    for (uint32_t j = h & (uint32_t)(cap - 1);; j = (j + 1) & (uint32_t)(cap - 1)) {
      uint32_t* e = &slots[(size_t)j * 4];
      if (e[3] == 0xffffffffu) {
        e[0] = (uint32_t)first8;
        e[1] = (uint32_t)(first8 >> 32);
        e[2] = h;
        e[3] = (uint32_t)l.size() | (uint32_t)r << 16;
        break;
      }
      if (f->labels[e[3] >> 16] == l) return fail("duplicate label '%s'", l.c_str());
    }
  }
  for (int r = 0; r < T; ++r) f->label_ptrs[r] = f->labels[r].c_str();
  chars.resize((chars.size() + 16) / 16 * 16, 0);
  if (f->d_label_slots.reserve(slots.size() * 4) || f->d_label_off.reserve(off.size() * 4) || f->d_label_chars.reserve(chars.size()))
    return 1;
  // (ordered on the compute stream and waited for, like the tip codes: see configure_batch)
  LH_CUDA(cudaStreamSynchronize(f->stream));
  LH_CUDA(cudaStreamSynchronize(f->plan_stream));
  LH_CUDA(cudaMemcpyAsync(f->d_label_slots.p, slots.data(), slots.size() * 4, cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaMemcpyAsync(f->d_label_off.p, off.data(), off.size() * 4, cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaMemcpyAsync(f->d_label_chars.p, chars.data(), chars.size(), cudaMemcpyHostToDevice, f->stream));
  LH_CUDA(cudaStreamSynchronize(f->stream));
  f->lt.slots = f->d_label_slots.as<uint4>();
  f->lt.mask = (uint32_t)(cap - 1);
  f->lt.long_off = f->d_label_off.as<uint32_t>();
  f->lt.chars = f->d_label_chars.as<char>();
  f->labels_set = true;
  return 0;
}

int lh_set_option(lh_family_t* f, const char* key, int64_t value) {
  std::string k(key ? key : "");
  if (k == "sites_per_thread") {
    if (value < 0 || value > 4) return fail("sites_per_thread must be 0 (auto) or 1..4");
    f->opt_sites_per_thread = (int)value;
  } else if (k == "smem_slots") {
    f->opt_smem_slots = (int)value;
  } else if (k == "smem_lo") {
    f->opt_smem_lo = (int)value;
  } else if (k == "site_patterns") {
    f->opt_site_patterns = value != 0;
  } else if (k == "chunk_growth") {
    f->opt_chunk_growth = value != 0;
  } else if (k == "open_unit") {
    f->opt_open_unit = value != 0;
  } else if (k == "parse_reserve") {
    if (value < 0 || value > 3) return fail("parse_reserve must be 0..3");
    f->opt_parse_reserve = (int)value;
  } else if (k == "keep_tables") {
    if (value != 0 && value != 1) return fail("keep_tables must be 0 or 1");
    f->opt_keep_tables = (int)value;
  } else if (k == "text_direct") {
    if (value != 0 && value != 1) return fail("text_direct must be 0 or 1");
    f->opt_text_direct = (int)value;
  } else if (k == "asr_scratch_mb") {
    if (value < 0) return fail("asr_scratch_mb must be >= 0 (0 = default, 4096)");
    f->opt_asr_scratch_mb = (int)value;
  } else if (k == "newick_kernel") {
    if (value < 0 || value > 3) return fail("newick_kernel must be 0 (default) / 2 (one CTA per tree), 1 (one thread per tree) or 3 (one warp per tree)");
    f->opt_newick_kernel = (int)value;
  } else if (k == "device_plan") {
    if (value < -1 || value > 1) return fail("device_plan must be -1 (auto), 0 or 1");
    f->opt_device_plan = (int)value;
  } else if (k == "hmm_overlap") {
    f->opt_hmm_overlap = value != 0;
  } else if (k == "timing") {
    f->opt_timing = value != 0;
  } else if (k == "host_threads") {
    f->opt_threads = (int)value;
    f->pool.reset();
  } else if (k == "chunk_trees") {
    if (value < 0) return fail("chunk_trees must be >= 0 (0 = auto)");
    f->opt_chunk_trees = (int)value;
  } else if (k == "pipe_waves") {
    if (value < 1) return fail("pipe_waves must be >= 1");
    f->opt_pipe_waves = (int)value;
  } else if (k == "chunk_bytes") {
    f->opt_chunk_bytes = (size_t)value;
  } else if (k == "stream") {
    // run on a caller-owned CUDA stream (e.g. torch.cuda.current_stream().cuda_stream); the value is the
    // cudaStream_t itself, so 0 is the legacy default stream (what torch's default stream is) and the
    // kernels are then ordered with everything the caller queues there; -1 restores a private stream
    cudaSetDevice(f->device);
    cudaStreamSynchronize(f->stream);
    if (f->stream && f->own_stream) cudaStreamDestroy(f->stream);
    if (value == -1) {
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

// Plans trees [t0, t0+n) into `steps` / `pblen` (segment-local, multi-threaded); raises
// *max_slots to the deepest register stack met.
int plan_range(lh_family* f, const lh_tree_batch_t* tb, int t0, int n, lhb::Step* steps, double* pblen,
               int* max_slots) {
  const int T = f->fd.T, n_ops = T - 2, NB = 2 * T - 3;
  if (!f->pool) {
    int nthreads = f->opt_threads > 0 ? f->opt_threads : (int)std::thread::hardware_concurrency();
    nthreads = std::max(1, std::min(nthreads, 32));
    f->pool.reset(new lhb::WorkerPool(nthreads - 1));  // the calling thread works too
    f->plan_scratch.resize(nthreads);
  }
  std::atomic<int> bad(-1), bad_len(-1), deepest(0);
  constexpr int kBlock = 16;  // trees per task
  const int n_tasks = (n + kBlock - 1) / kBlock;
  f->pool->run(n_tasks, [&](int task, int worker) {
    lhb::PlanScratch& sc = f->plan_scratch[worker];
    int mx = 0;
    const int lo = task * kBlock, hi = std::min(n, lo + kBlock);
    for (int i = lo; i < hi; ++i) {
      const int t = t0 + i;
      const double* bl = tb->blen + (size_t)t * n_ops * 2;
      bool neg = !(tb->root_blen[t] >= 0.0);
      for (int j = 0; j < 2 * n_ops; ++j) neg |= !(bl[j] >= 0.0);
      if (neg) {
        bad_len.store(t);
        return;
      }
      int s = lhb::plan_tree(T, tb->children + (size_t)t * n_ops * 2, bl, tb->root_child[t], tb->root_blen[t],
                             steps + (size_t)i * (T - 1), pblen + (size_t)i * NB, sc);
      if (s < 0) {
        bad.store(t);
        return;
      }
      mx = std::max(mx, s);
    }
    int cur = deepest.load();
    while (mx > cur && !deepest.compare_exchange_weak(cur, mx)) {}
  });
  if (bad_len.load() >= 0) return fail("tree %d: negative branch length", bad_len.load());
  if (bad.load() >= 0) return fail("tree %d is not a valid post-order operation list", bad.load());
  *max_slots = std::max(*max_slots, deepest.load());
  return 0;
}

int check_batch(lh_family* f, const lh_tree_batch_t* tb) {
  const int B = tb->n_trees, R = tb->n_rates, NB = 2 * f->fd.T - 3;
  if (B < 0) return fail("negative batch size");
  if (R < 1 || R > LH_MAX_RATES) return fail("n_rates %d outside 1..%d", R, LH_MAX_RATES);
  for (int t = 0; t < B; ++t)
    for (int i = 0; i < 4; ++i)
      if (!(tb->pi[4 * t + i] > 0.0)) return fail("tree %d: pi[%d] must be positive", t, i);
  // (branch lengths are checked by the planner threads, tree by tree, as they stream them)
  // A tree can run on site patterns when the naive tip (xMSA row 0) is on the root edge or under
  // the virtual root (the last operation): then no other node's CLV depends on the naive row.
  const int n_ops = f->fd.T - 2;
  f->on_patterns.assign((size_t)B, 0);
  f->n_pattern_trees = 0;
  if (f->opt_site_patterns && f->cols[1].n_cols < f->cols[0].n_cols)
    for (int t = 0; t < B; ++t) {
      const int32_t* last = tb->children + ((size_t)t * n_ops + (n_ops - 1)) * 2;
      const bool ok = tb->root_child[t] == 0 || last[0] == 0 || last[1] == 0;
      f->on_patterns[t] = ok;
      f->n_pattern_trees += ok;
    }
  return 0;
}

// Device buffers and launch configuration for a batch of B trees with R rate categories.
// Everything here depends on (family, B, R) only, so it can run before any tree is planned: the
// number of CLV slots is bounded by the Strahler bound floor(log2 T) + 1 of a T-tip tree rather
// than by the planned maximum.
int configure_batch(lh_family* f, int B, int R) {
  const int T = f->fd.T, S = f->fd.S, NB = 2 * T - 3;
  f->B = B;
  f->R = R;
  f->slots = 0;
  f->window_set = false;
  f->tm = lh_timings_t{};
  f->tm.n_trees = B;
  if (B == 0) return 0;
  const size_t plan_bytes = (size_t)B * (T - 1) * sizeof(lhb::Step);
  const size_t pblen_bytes = (size_t)B * NB * sizeof(double);
  const size_t misc_doubles = (size_t)B * (6 + 4 + R);
  if (f->d_plan.reserve(plan_bytes) || f->d_pblen.reserve(pblen_bytes) ||
      f->d_er.reserve(misc_doubles * sizeof(double)) || f->d_eig.reserve((size_t)B * lhb::kEigStride * 8) ||
      f->d_xe.reserve((size_t)B * S * 8) || f->d_loglik.reserve((size_t)B * 8) ||
      f->h_misc.reserve(misc_doubles * sizeof(double)))
    return 1;
  {
    // Who plans the traversals. The host planner threads cost the GPU nothing as long as they stay
    // ahead of the kernels (about 20 ns per operation and thread against 2.7 ps per operation, column
    // and category of k_prune, both measured on C4); k_plan costs about 2 % of a step but no host time.
    // Auto: the device plans when the host threads this handle has would need more than half the
    // kernel time (few threads per GPU: 8 ranks on one host, repertoire slots).
    int threads = f->opt_threads > 0 ? f->opt_threads : (int)std::thread::hardware_concurrency();
    threads = std::max(1, std::min(threads, 32));
    const int cols = (f->opt_site_patterns && f->n_pattern_trees * 2 >= B) ? f->cols[1].n_cols : f->cols[0].n_cols;
    const double host_ns = 20.0 / threads, kernel_ns = 2.7e-3 * cols * R;
    // A batch of one or two chunks has nothing to hide its first chunk's planning behind: there the
    // device planner's shorter lead-in wins whatever the thread count (Liao, 1024 samples, 16 threads:
    // 3.48 ms per call with the host planner, 2.45 ms with k_plan).
    const bool small = (size_t)B * lhb::records_per_pass(T) * 20 * R * sizeof(double) <= ((size_t)1 << 30);
    const bool want = f->opt_device_plan < 0 ? (small || host_ns > 0.5 * kernel_ns) : f->opt_device_plan != 0;
    f->dev_plan = want && lhb::plan_on_device_ok(T);
    if (f->newick_mode) {  // k_newick feeds k_plan
      if (!lhb::plan_on_device_ok(T)) return fail("the device Newick parser takes families of up to 32769 tips (got %d)", T);
      f->dev_plan = true;
    }
    // how much faster the host stages a tree than the GPU scores it (the device planner leaves the host
    // a memcpy: ~6 ns per operation and thread)
    f->stage_ratio = kernel_ns / (f->dev_plan ? 6.0 / threads : host_ns);
  }
  f->flags_pending = false;
  f->run_recorded = false;
  if (f->newick_mode) {
    if (f->d_on_pat.reserve((size_t)B) || f->d_nw_status.reserve((size_t)B) || f->h_nw_status.reserve((size_t)B)) return 1;
    if (f->d_toffs.reserve((size_t)B * 8) || f->d_tlens.reserve((size_t)B * 4)) return 1;
  }
  if (f->dev_plan) {
    const size_t n2 = (size_t)B * (T - 2) * 2;
    if (f->d_rchildren.reserve(n2 * sizeof(int32_t)) || f->d_rblen.reserve(n2 * sizeof(double)) ||
        f->d_rroot_child.reserve((size_t)B * sizeof(int32_t)) || f->d_rroot_blen.reserve((size_t)B * sizeof(double)) ||
        f->d_flags.reserve(4 * sizeof(int32_t)) || f->h_flags.reserve(8 * sizeof(int32_t)))
      return 1;
  }
  lhb::BatchDev& bd = f->bd;
  bd = lhb::BatchDev{};
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

  // pruning launch configuration per column set in use: [0] all columns, [1] site patterns
  const size_t rec_bytes_tree = (size_t)lhb::records_per_pass(T) * 20 * R * sizeof(double);
  f->rec_bytes_tree = rec_bytes_tree;
  int strahler = 1;
  while ((2 << (strahler - 1)) <= T) ++strahler;  // floor(log2 T) + 1
  // (text input: which trees run on patterns is only known on the device, so both sets are made ready)
  const bool use[2] = {f->n_pattern_trees < B || f->newick_mode, f->n_pattern_trees > 0};
  size_t scratch_bytes = 0, scratch_cnt_bytes = 0;
  for (int g = 0; g < 2; ++g) {
    lhb::PruneCfg& pc = f->pc[g];
    pc = lhb::PruneCfg{};
    if (!use[g]) continue;
    lh_family::ColumnSet& cset = f->cols[g];
    const int NC = cset.n_cols;
    pc.n_cols = NC;
    pc.defer_naive = g;
    pc.sites_per_thread = f->opt_sites_per_thread ? f->opt_sites_per_thread : pick_sites_per_thread(NC, 32);
    const int tile_sites = 32 * pc.sites_per_thread;
    pc.n_tiles = (NC + tile_sites - 1) / tile_sites;
    pc.cols_pad = pc.n_tiles * tile_sites;
    pc.total_slots = strahler + 1;  // + the virtual-root edge
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
    f->occ[g] = lhb::prune_max_ctas_per_sm(pc.sites_per_thread, pc.smem_slots);
    if (f->occ[g] < 1) return fail("pruning kernel cannot be resident with %d shared-memory slots", pc.smem_slots);
    // Text input: the parser (k_newick_cta, 128 threads) of the next chunks runs beside the persistent pruning
    // CTAs, which otherwise hold every register of an SM until their launch ends: leave it room.
    if (f->newick_mode && f->occ[g] >= 4) f->occ[g] -= f->opt_parse_reserve;
    pc.log_threshold = std::log(1.0 / std::pow(2.0, 256));
    if (cset.codes_ns != pc.sites_per_thread) {
      // re-tile the tip codes for this NS (once per family, column set and NS)
      const int NS = pc.sites_per_thread;
      if ((uint64_t)T * pc.n_tiles * 32 >= (1ull << 32)) return fail("tip-code table too large");  // the kernel indexes it with 32 bits
      std::vector<uint32_t> ct((size_t)T * pc.n_tiles * 32, 0u);
      for (int r = 0; r < T; ++r) {
        const uint8_t* row = &cset.codes[(size_t)r * NC];
        for (int tile = 0; tile < pc.n_tiles; ++tile) {
          uint32_t* w = &ct[((size_t)r * pc.n_tiles + tile) * 32];
          for (int u = 0; u < NS; ++u) {
            const int s0 = tile * 32 * NS + u * 32, n = std::min(32, NC - s0);
            for (int lane = 0; lane < n; ++lane) w[lane] |= (uint32_t)(row[s0 + lane] * 32u) << (8 * u);
          }
        }
      }
      // The copy is ordered on the compute stream, in front of the kernels that read the table, and is
      // waited for. (A synchronous cudaMemcpy from pageable memory returns once the data is in the
      // driver's staging buffer; its DMA runs on the legacy stream, with which these non-blocking streams
      // do not synchronise - with several handles keeping the copy engine busy, k_prune of a rebound
      // handle could start on the previous family's codes.)
      LH_CUDA(cudaStreamSynchronize(f->stream));
      if (cset.d_codes_t.reserve(ct.size() * 4) || cset.h_codes_t.reserve(ct.size() * 4)) return 1;
      stream_copy(cset.h_codes_t.p, ct.data(), ct.size() * 4);
      LH_CUDA(cudaMemcpyAsync(cset.d_codes_t.p, cset.h_codes_t.p, ct.size() * 4, cudaMemcpyHostToDevice, f->stream));
      LH_CUDA(cudaStreamSynchronize(f->stream));
      cset.codes_ns = NS;
    }
    pc.codes_t = cset.d_codes_t.as<uint32_t>();
  }

  // chunk the batch: one launch of each kernel per chunk (k_prune: one per column set in use).
  // The chunk is sized on the column set most trees use.
  const int gmain = f->n_pattern_trees * 2 >= B ? 1 : 0;
  // A wave = the trees whose work items fill the resident CTAs exactly once. CTAs stride over the
  // items of a launch and items of one family cost the same, so a chunk of whole waves leaves no
  // partial last wave.
  auto gcd = [](int a, int b) { while (b) { int r = a % b; a = b; b = r; } return a; };
  // (a work item of k_prune is one (tree, tile, category))
  const int resident = f->occ[gmain] * f->num_sms, tiles = f->pc[gmain].n_tiles, items = tiles * R;
  const int wave = resident / gcd(resident, items);  // smallest tree count with trees*items % resident == 0
  int chunk = (int)std::max<size_t>(1, f->opt_chunk_bytes / rec_bytes_tree);
  // A resident batch runs 16 waves per launch (the records then exceed L2 for large trees, which is
  // harmless: they are written once and read once per site tile by CTAs that run at the same time,
  // ~2.6 MB of DRAM traffic per 1001-tip tree).
  const int waves16 = (16 * resident + tiles - 1) / tiles;
  chunk = std::max(chunk, waves16);
  if (chunk > wave) chunk -= chunk % wave;
  // The pipelined call launches every `pipe_waves` waves, so its first kernels start after a
  // small part of the host planning and the copy of one chunk hides behind the kernels of another.
  f->wave_unit = wave;
  f->wave_pipe = (std::max(wave, (resident + tiles - 1) / tiles) + wave - 1) / wave * wave;
  int chunk_pipe = std::max(1, f->opt_pipe_waves) * f->wave_pipe;
  if (f->opt_chunk_trees > 0) {
    chunk = chunk_pipe = f->opt_chunk_trees;
    f->wave_pipe = 0;
  }
  chunk = std::min(chunk, B);
  f->chunk = chunk;
  f->chunk_pipe = std::min(chunk_pipe, chunk);
  if (f->d_prec.reserve((size_t)chunk * rec_bytes_tree)) return 1;
  bd.prec = f->d_prec.as<double>();
  const int pad_max = std::max(f->pc[0].cols_pad, f->pc[1].cols_pad);
  const int64_t e_stride = (int64_t)R * 8 * pad_max, c_stride = (int64_t)R * pad_max;
  if (f->d_ebuf.reserve((size_t)chunk * e_stride * sizeof(double)) ||
      f->d_ecnt.reserve((size_t)chunk * c_stride * sizeof(int32_t)) ||
      f->d_enaive.reserve((size_t)chunk * sizeof(int32_t)) || f->d_list.reserve((size_t)B * sizeof(int32_t)))
    return 1;
  for (int g = 0; g < 2; ++g) {
    lhb::PruneCfg& pc = f->pc[g];
    if (!use[g]) continue;
    const int tile_sites = 32 * pc.sites_per_thread;
    pc.grid = (int)std::min<int64_t>((int64_t)f->occ[g] * f->num_sms, (int64_t)chunk * pc.n_tiles * R);
    const int gslots = pc.total_slots - pc.smem_slots;
    if (gslots > 0) {
      scratch_bytes = std::max(scratch_bytes, (size_t)pc.grid * gslots * 4 * tile_sites * sizeof(double));
      scratch_cnt_bytes = std::max(scratch_cnt_bytes, (size_t)pc.grid * gslots * tile_sites * sizeof(int));
    }
    pc.ebuf = f->d_ebuf.as<double>();
    pc.ecnt = f->d_ecnt.as<int32_t>();
    pc.enaive = f->d_enaive.as<int32_t>();
    pc.e_stride = e_stride;
    pc.c_stride = c_stride;
  }
  if (f->d_scratch.reserve(scratch_bytes) || f->d_scratch_cnt.reserve(scratch_cnt_bytes)) return 1;
  for (auto& pc : f->pc) {
    pc.scratch = f->d_scratch.as<double>();
    pc.scratch_cnt = f->d_scratch_cnt.as<int32_t>();
  }
  lhb::EdgeCfg& ec = f->ec;
  ec = lhb::EdgeCfg{};
  ec.ebuf = f->d_ebuf.as<double>();
  ec.ecnt = f->d_ecnt.as<int32_t>();
  ec.enaive = f->d_enaive.as<int32_t>();
  ec.e_stride = e_stride;
  ec.c_stride = c_stride;
  ec.cols_pad_all = f->pc[0].cols_pad;
  ec.cols_pad_pat = f->pc[1].cols_pad;
  ec.col_of_site = f->d_col_of_site;
  ec.log_threshold = std::log(1.0 / std::pow(2.0, 256));
  if (f->want_rate_ll) {
    if (f->d_rate_ll.reserve((size_t)B * R * S * sizeof(double))) return 1;
    ec.rate_ll = f->d_rate_ll.as<double>();
  }

  // staging segments: up to ~48 MB of plan + branch lengths per ring slot, never more than a
  // pipelined chunk
  const size_t stage_bytes_tree = (size_t)(T - 1) * sizeof(lhb::Step) + (size_t)NB * sizeof(double);
  int seg = (int)std::max<size_t>(64, ((size_t)48 << 20) / stage_bytes_tree);
  f->seg = std::min(seg, f->chunk_pipe);
  for (auto& r : f->h_ring)
    if (r.reserve(f->newick_mode ? kTextSlotBytes : (size_t)f->seg * stage_bytes_tree)) return 1;

  lhb::HmmOut& ho = f->ho;
  ho = lhb::HmmOut{};
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
  const lhb::PruneCfg& pm = f->pc[gmain];
  f->tm.prune_slots = pm.total_slots;
  f->tm.prune_tile_sites = 32 * pm.sites_per_thread;
  f->tm.prune_sites_per_thread = pm.sites_per_thread;
  f->tm.prune_smem_slots = pm.smem_slots;
  f->tm.prune_grid = pm.grid;
  f->tm.prune_ctas_per_sm = f->occ[gmain];
  f->tm.prune_columns = pm.n_cols;
  f->tm.pattern_trees = f->n_pattern_trees;
  f->tm.device_plan = f->dev_plan ? 1 : 0;
  return 0;
}

void launch_plan_trees(lh_family* f, int t0, int n, cudaStream_t s);

// Plans trees [t0, t0+n) into ring slot `slot` and queues their copy to HBM on `s`.
int stage_segment(lh_family* f, const lh_tree_batch_t* tb, int t0, int n, int slot, cudaStream_t s) {
  const int T = f->fd.T, NB = 2 * T - 3;
  if (f->ring_busy[slot]) {
    LH_CUDA(cudaEventSynchronize(f->ring_ev[slot]));
    f->ring_busy[slot] = false;
  }
  lhb::Step* steps = f->h_ring[slot].as<lhb::Step>();
  double* pblen = reinterpret_cast<double*>(steps + (size_t)n * (T - 1));
  const auto t_a = std::chrono::steady_clock::now();
  // Device planning: the segment's operation lists are only copied into the pinned slot (the host
  // threads share the memcpy); the first trees of a batch are still planned here, into a scratch,
  // because their steps choose the shared-memory window of the CLV stack.
  std::vector<lhb::Step> probe_steps;
  std::vector<double> probe_pblen;
  const int n_planned = f->dev_plan ? (f->window_set ? 0 : std::min(n, 64)) : n;
  if (f->dev_plan && n_planned) {
    probe_steps.resize((size_t)n_planned * (T - 1));
    probe_pblen.resize((size_t)n_planned * NB);
    steps = probe_steps.data();
    pblen = probe_pblen.data();
  }
  if (n_planned && plan_range(f, tb, t0, n_planned, steps, pblen, &f->slots)) return 1;
  if (!f->window_set) {
    // which CLV stack levels live in shared memory: the window of smem_slots consecutive levels
    // that the first trees of the batch park results in most often
    int64_t hist[32] = {};
    const int sample = std::min(n, 64);
    for (int i = 0; i < sample; ++i)
      for (int q = 0; q < T - 2; ++q) {
        const int dst = steps[(size_t)i * (T - 1) + q].dst;
        if (dst >= 0 && dst < 32) ++hist[dst];
      }
    for (auto& pc : f->pc) {
      int best = 0;
      int64_t best_sum = -1;
      for (int lo = 0; lo + pc.smem_slots <= std::max(pc.total_slots, pc.smem_slots); ++lo) {
        int64_t sum = 0;
        for (int k = lo; k < lo + pc.smem_slots && k < 32; ++k) sum += hist[k];
        if (sum > best_sum) { best_sum = sum; best = lo; }
      }
      pc.smem_lo = f->opt_smem_lo >= 0 ? std::min(f->opt_smem_lo, std::max(0, pc.total_slots - pc.smem_slots)) : best;
    }
    f->tm.prune_smem_lo = f->pc[f->n_pattern_trees * 2 >= f->B ? 1 : 0].smem_lo;
    f->window_set = true;
  }
  f->tm.ms_plan += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_a).count();
  const int slot_bound = std::max(f->pc[0].total_slots, f->pc[1].total_slots);
  if (f->slots > slot_bound) return fail("internal: tree needs %d CLV slots, bound is %d", f->slots, slot_bound);
  if (f->dev_plan) {
    // raw layout of a slot: children [n][T-2][2] | blen [n][T-2][2] | root_blen [n] | root_child [n]
    const auto t_b = std::chrono::steady_clock::now();
    const size_t per = (size_t)(T - 2) * 2;
    uint8_t* base = f->h_ring[slot].as<uint8_t>();
    int32_t* r_ch = reinterpret_cast<int32_t*>(base);
    double* r_bl = reinterpret_cast<double*>(base + ((size_t)n * per * sizeof(int32_t) + 63) / 64 * 64);
    double* r_rb = r_bl + (size_t)n * per;
    int32_t* r_rc = reinterpret_cast<int32_t*>(r_rb + n);
    if (!f->pool) {
      int nthreads = f->opt_threads > 0 ? f->opt_threads : (int)std::thread::hardware_concurrency();
      nthreads = std::max(1, std::min(nthreads, 32));
      f->pool.reset(new lhb::WorkerPool(nthreads - 1));
      f->plan_scratch.resize(nthreads);
    }
    constexpr int kBlock = 32;  // trees per task
    f->pool->run((n + kBlock - 1) / kBlock, [&](int task, int) {
      const int lo = task * kBlock, cnt = std::min(n, lo + kBlock) - lo;
      stream_copy(r_ch + (size_t)lo * per, tb->children + ((size_t)t0 + lo) * per, (size_t)cnt * per * sizeof(int32_t));
      stream_copy(r_bl + (size_t)lo * per, tb->blen + ((size_t)t0 + lo) * per, (size_t)cnt * per * sizeof(double));
    });
    std::memcpy(r_rb, tb->root_blen + t0, (size_t)n * sizeof(double));
    std::memcpy(r_rc, tb->root_child + t0, (size_t)n * sizeof(int32_t));
    f->tm.ms_plan += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_b).count();
    LH_CUDA(cudaMemcpyAsync(f->d_rchildren.as<int32_t>() + (size_t)t0 * per, r_ch, (size_t)n * per * sizeof(int32_t),
                            cudaMemcpyHostToDevice, s));
    LH_CUDA(cudaMemcpyAsync(f->d_rblen.as<double>() + (size_t)t0 * per, r_bl, (size_t)n * per * sizeof(double),
                            cudaMemcpyHostToDevice, s));
    LH_CUDA(cudaMemcpyAsync(f->d_rroot_blen.as<double>() + t0, r_rb, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    LH_CUDA(cudaMemcpyAsync(f->d_rroot_child.as<int32_t>() + t0, r_rc, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    LH_CUDA(cudaEventRecord(f->ring_ev[slot], s));
    f->ring_busy[slot] = true;
    f->tm.h2d_bytes += (int64_t)((size_t)n * per * 12 + (size_t)n * 12);
    return 0;
  }
  const size_t plan_bytes = (size_t)n * (T - 1) * sizeof(lhb::Step), pblen_bytes = (size_t)n * NB * sizeof(double);
  LH_CUDA(cudaMemcpyAsync(f->d_plan.as<lhb::Step>() + (size_t)t0 * (T - 1), steps, plan_bytes,
                          cudaMemcpyHostToDevice, s));
  LH_CUDA(cudaMemcpyAsync(f->d_pblen.as<double>() + (size_t)t0 * NB, pblen, pblen_bytes, cudaMemcpyHostToDevice, s));
  LH_CUDA(cudaEventRecord(f->ring_ev[slot], s));
  f->ring_busy[slot] = true;
  f->tm.h2d_bytes += (int64_t)(plan_bytes + pblen_bytes);
  return 0;
}

void ensure_pool(lh_family* f) {
  if (f->pool) return;
  int nthreads = f->opt_threads > 0 ? f->opt_threads : (int)std::thread::hardware_concurrency();
  nthreads = std::max(1, std::min(nthreads, 32));
  f->pool.reset(new lhb::WorkerPool(nthreads - 1));  // the calling thread works too
  f->plan_scratch.resize(nthreads);
}

// Text input, first trees of a batch: parsed and planned on the host, into a scratch, because their
// programs choose the shared-memory window of the CLV stack (the same probe as for array input).
int probe_window_from_text(lh_family* f, const lh_newick_batch_t* nb, int t0, int n) {
  const int T = f->fd.T, NB = 2 * T - 3;
  const size_t per = (size_t)(T - 2) * 2;
  std::vector<int32_t> ch(n * per), rc(n);
  std::vector<double> bl(n * per), rb(n);
  if (lhh_newick_to_ops(T, f->label_ptrs.data(), n, nb->newick + t0, f->opt_threads, ch.data(), bl.data(), rc.data(), rb.data()))
    return 0;  // a malformed tree among them: the device pass flags it and the call reports it; any window is valid
  lh_tree_batch_t tb{};
  tb.n_trees = n;
  tb.children = ch.data();
  tb.blen = bl.data();
  tb.root_child = rc.data();
  tb.root_blen = rb.data();
  std::vector<lhb::Step> steps((size_t)n * (T - 1));
  std::vector<double> pblen((size_t)n * NB);
  int slots = 0;
  ensure_pool(f);
  {
    // (plan_range reports bad trees through fail(); here they are left to the device pass)
    std::atomic<int> bad(0);
    lhb::PlanScratch& sc = f->plan_scratch[f->pool->workers()];
    for (int i = 0; i < n; ++i) {
      const int d = lhb::plan_tree(T, ch.data() + i * per, bl.data() + i * per, rc[i], rb[i], steps.data() + (size_t)i * (T - 1),
                                   pblen.data() + (size_t)i * NB, sc);
      if (d < 0) return 0;
      slots = std::max(slots, d);
    }
  }
  f->slots = std::max(f->slots, slots);
  int64_t hist[32] = {};
  for (int i = 0; i < n; ++i)
    for (int q = 0; q < T - 2; ++q) {
      const int dst = steps[(size_t)i * (T - 1) + q].dst;
      if (dst >= 0 && dst < 32) ++hist[dst];
    }
  for (auto& pc : f->pc) {
    int best = 0;
    int64_t best_sum = -1;
    for (int lo = 0; lo + pc.smem_slots <= std::max(pc.total_slots, pc.smem_slots); ++lo) {
      int64_t sum = 0;
      for (int k = lo; k < lo + pc.smem_slots && k < 32; ++k) sum += hist[k];
      if (sum > best_sum) { best_sum = sum; best = lo; }
    }
    pc.smem_lo = f->opt_smem_lo >= 0 ? std::min(f->opt_smem_lo, std::max(0, pc.total_slots - pc.smem_slots)) : best;
  }
  f->tm.prune_smem_lo = f->pc[1].n_cols ? f->pc[1].smem_lo : f->pc[0].smem_lo;
  return 0;
}

// Copies the strings of trees [t0, t0+n) (lengths len[0..n)) into ring slot `slot` and queues their copy on `cs`
// into the chunk's device text buffer at byte `text_at` (returns the bytes used through *used).
// The shared-memory window of the CLV stack for a text call, from the shapes of the first trees (once per call).
int choose_window_from_text(lh_family* f, const lh_newick_batch_t* nb, int t0, int n) {
  if (!f->window_set) {
    // The window is a placement choice (results do not depend on it) made from the shapes of the first trees.
    // Parsing them on the host costs 0.16 ms per 1001-tip string, serially, in front of the call's first kernel:
    // the choice is kept for the bound family while the stack geometry is the same, and the probe is short.
    lh_family::WindowCache& wc = f->win_cache;
    const bool hit = wc.valid && wc.opt_lo == f->opt_smem_lo && wc.smem_slots[0] == f->pc[0].smem_slots &&
                     wc.smem_slots[1] == f->pc[1].smem_slots && wc.total_slots[0] == f->pc[0].total_slots &&
                     wc.total_slots[1] == f->pc[1].total_slots;
    if (hit) {
      f->pc[0].smem_lo = wc.lo[0];
      f->pc[1].smem_lo = wc.lo[1];
      f->slots = std::max(f->slots, wc.slots);
      f->tm.prune_smem_lo = f->pc[1].n_cols ? f->pc[1].smem_lo : f->pc[0].smem_lo;
    } else {
      if (probe_window_from_text(f, nb, t0, std::min(n, 8))) return 1;
      wc.valid = true;
      wc.opt_lo = f->opt_smem_lo;
      for (int g = 0; g < 2; ++g) {
        wc.lo[g] = f->pc[g].smem_lo;
        wc.smem_slots[g] = f->pc[g].smem_slots;
        wc.total_slots[g] = f->pc[g].total_slots;
      }
      wc.slots = f->slots;
    }
    f->window_set = true;
  }
  return 0;
}

// Slot layout: offs [n] uint64 (into the chunk's text buffer) | lens [n] uint32 | pad to 16 | strings, each
// NUL-terminated and 16-byte aligned.
int stage_text_segment(lh_family* f, const lh_newick_batch_t* nb, int t0, int n, const int64_t* len, int slot,
                       char* d_chunk_text, size_t text_at, size_t* used, cudaStream_t cs) {
  if (f->ring_busy[slot]) {
    LH_CUDA(cudaEventSynchronize(f->ring_ev[slot]));  // the slot's previous segment has been copied
    f->ring_busy[slot] = false;
  }
  const auto t_a = std::chrono::steady_clock::now();
  if (choose_window_from_text(f, nb, t0, n)) return 1;
  uint8_t* base = f->h_ring[slot].as<uint8_t>();
  uint64_t* offs = reinterpret_cast<uint64_t*>(base);
  uint32_t* lens32 = reinterpret_cast<uint32_t*>(offs + n);
  const size_t text0 = ((size_t)n * 12 + 15) / 16 * 16;
  size_t at = 0;
  for (int i = 0; i < n; ++i) {
    offs[i] = (uint64_t)(text_at + at);
    lens32[i] = (uint32_t)len[i];
    at += ((size_t)len[i] + 1 + 15) / 16 * 16;
  }
  ensure_pool(f);
  constexpr int kBlock = 8;  // trees per task
  f->pool->run((n + kBlock - 1) / kBlock, [&](int task, int) {
    const int lo = task * kBlock, hi = std::min(n, lo + kBlock);
    for (int i = lo; i < hi; ++i) {
      uint8_t* dst = base + text0 + (size_t)(offs[i] - text_at);
      const size_t l = (size_t)len[i];
      const size_t body = l / 16 * 16;  // whole 16-byte blocks stream straight from the caller's string
      stream_copy(dst, nb->newick[t0 + i], body);
      uint8_t tail[16] = {};
      std::memcpy(tail, nb->newick[t0 + i] + body, l - body);   // the NUL and the padding come with it
      stream_copy(dst + body, tail, 16);
    }
  });
  _mm_sfence();
  f->tm.ms_plan += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_a).count();
  LH_CUDA(cudaMemcpyAsync(d_chunk_text + text_at, base + text0, at, cudaMemcpyHostToDevice, cs));
  LH_CUDA(cudaMemcpyAsync(f->d_toffs.as<uint64_t>() + t0, offs, (size_t)n * 8, cudaMemcpyHostToDevice, cs));
  LH_CUDA(cudaMemcpyAsync(f->d_tlens.as<uint32_t>() + t0, lens32, (size_t)n * 4, cudaMemcpyHostToDevice, cs));
  LH_CUDA(cudaEventRecord(f->ring_ev[slot], cs));
  f->ring_busy[slot] = true;
  f->tm.h2d_bytes += (int64_t)(at + (size_t)n * 12);
  *used = at;
  return 0;
}

// The strings of the chunk's trees [c0, c0+cn) -> one device text buffer (two alternate between chunks), in
// segments of up to kTextSlotBytes through the pinned ring, then ONE parse of the whole chunk on the plan
// stream: a launch of a few hundred trees leaves most of the GPU idle, and the parser of a chunk can only run
// in the gaps the persistent pruning CTAs of the chunk before leave anyway. *k counts ring slots.
int stage_text_chunk(lh_family* f, const lh_newick_batch_t* nb, int ci, int c0, int cn, int* k, cudaStream_t cs) {
  std::vector<int64_t> lens(cn);
  if (nb->length) {
    std::copy(nb->length + c0, nb->length + c0 + cn, lens.begin());
  } else {
    ensure_pool(f);
    f->pool->run((cn + 15) / 16, [&](int task, int) {
      for (int i = task * 16; i < std::min(cn, task * 16 + 16); ++i) lens[i] = (int64_t)std::strlen(nb->newick[c0 + i]);
    });
  }
  size_t chunk_bytes = 0;
  for (int i = 0; i < cn; ++i) {
    if (lens[i] < 0 || (uint64_t)lens[i] + 64 > kTextSlotBytes)
      return fail("tree %d: a Newick string of %lld bytes is too long", c0 + i, (long long)lens[i]);
    chunk_bytes += ((size_t)lens[i] + 1 + 15) / 16 * 16;
  }
  // Text the caller holds in pinned memory at a constant stride (one buffer, a string every `stride` bytes - how a
  // reader that knows the longest row lays them out, and how host.NewickBuffer does) is copied in place: one 2-D
  // copy per chunk, `width` = the chunk's longest string, no pass of the host over the bytes. Anything else goes
  // through the pinned ring below.
  bool direct = f->text_direct_call && ((uintptr_t)nb->newick[c0] & 7u) == 0;
  int64_t stride = 0;
  size_t width = 0;
  if (direct) {
    int64_t maxlen = 0;
    for (int i = 0; i < cn; ++i) maxlen = std::max(maxlen, lens[i]);
    width = ((size_t)maxlen + 1 + 15) / 16 * 16;
    stride = cn > 1 ? (int64_t)(nb->newick[c0 + 1] - nb->newick[c0]) : (int64_t)width;
    if (cn == 1) width = (size_t)maxlen + 1;   // (a single string: nothing may be read behind its NUL)
    direct = stride > 0 && (stride & 7) == 0 && (int64_t)width <= stride;
    for (int i = 0; direct && i + 1 < cn; ++i) direct = nb->newick[c0 + i + 1] - nb->newick[c0 + i] == stride;
    if (direct) chunk_bytes = (size_t)cn * ((width + 15) / 16 * 16);
  }
  DevBuf& buf = f->d_textbuf[ci & 1];
  cudaStream_t ps = f->plan_stream;
  // the buffer's previous chunk (ci - 2) must have been parsed before it is overwritten - or freed
  if (ci >= 2) {
    if (chunk_bytes > buf.cap) LH_CUDA(cudaEventSynchronize(f->parsed_ev[ci & 1]));
    else LH_CUDA(cudaStreamWaitEvent(cs, f->parsed_ev[ci & 1], 0));
  }
  if (buf.reserve(chunk_bytes + 64)) return 1;
  size_t text_at = 0;
  if (direct) {
    const int slot = *k % lh_family::kRing;
    if (f->ring_busy[slot]) {
      LH_CUDA(cudaEventSynchronize(f->ring_ev[slot]));
      f->ring_busy[slot] = false;
    }
    if (choose_window_from_text(f, nb, c0, cn)) return 1;
    const size_t dpitch = (width + 15) / 16 * 16;
    uint64_t* offs = f->h_ring[slot].as<uint64_t>();
    uint32_t* lens32 = reinterpret_cast<uint32_t*>(offs + cn);
    for (int i = 0; i < cn; ++i) {
      offs[i] = (uint64_t)i * dpitch;
      lens32[i] = (uint32_t)lens[i];
    }
    // (every row but the last is followed by another string, so `width` bytes of it lie inside the caller's buffer;
    // the last row is copied with its own length: the buffer may end right behind its NUL)
    if (cn > 1)
      LH_CUDA(cudaMemcpy2DAsync(buf.as<char>(), dpitch, nb->newick[c0], (size_t)stride, width, (size_t)(cn - 1),
                                cudaMemcpyHostToDevice, cs));
    LH_CUDA(cudaMemcpyAsync(buf.as<char>() + (size_t)(cn - 1) * dpitch, nb->newick[c0 + cn - 1], (size_t)lens[cn - 1] + 1,
                            cudaMemcpyHostToDevice, cs));
    LH_CUDA(cudaMemcpyAsync(f->d_toffs.as<uint64_t>() + c0, offs, (size_t)cn * 8, cudaMemcpyHostToDevice, cs));
    LH_CUDA(cudaMemcpyAsync(f->d_tlens.as<uint32_t>() + c0, lens32, (size_t)cn * 4, cudaMemcpyHostToDevice, cs));
    LH_CUDA(cudaEventRecord(f->ring_ev[slot], cs));
    f->ring_busy[slot] = true;
    f->tm.h2d_bytes += (int64_t)((size_t)cn * width + (size_t)cn * 12);
    ++f->tm.text_direct;
    ++*k;
  }
  for (int i = direct ? cn : 0; i < cn;) {
    size_t bytes = 0;
    int n = 0;
    while (i + n < cn) {
      const size_t add = ((size_t)lens[i + n] + 1 + 15) / 16 * 16 + 12;
      if (n && bytes + add + 16 > kTextSegBytes) break;
      bytes += add;
      ++n;
    }
    size_t used = 0;
    if (stage_text_segment(f, nb, c0 + i, n, lens.data() + i, *k % lh_family::kRing, buf.as<char>(), text_at, &used, cs)) return 1;
    text_at += used;
    i += n;
    ++*k;
  }
  LH_CUDA(cudaEventRecord(f->text_copied, cs));
  LH_CUDA(cudaStreamWaitEvent(ps, f->text_copied, 0));
  lhb::NewickDev p{};
  p.T = f->fd.T;
  p.tree0 = c0;
  p.n = cn;
  p.text = buf.as<char>();
  p.offs = f->d_toffs.as<uint64_t>() + c0;
  p.lens = f->d_tlens.as<uint32_t>() + c0;
  p.kernel = f->opt_newick_kernel;
  if (p.kernel != 1) {
    // (one scratch: the parses of a call run one after the other on the plan stream; growing it waits for them)
    const size_t need = lhb::newick_items_bytes(p.T, cn);
    if (need > f->d_items.cap) LH_CUDA(cudaStreamSynchronize(ps));
    if (f->d_items.reserve(need)) return 1;
    p.items = f->d_items.as<unsigned char>();
  }
  p.labels = f->lt;
  p.children = f->d_rchildren.as<int32_t>();
  p.blen = f->d_rblen.as<double>();
  p.root_child = f->d_rroot_child.as<int32_t>();
  p.root_blen = f->d_rroot_blen.as<double>();
  p.on_pat = f->d_on_pat.as<uint8_t>();
  p.status = f->d_nw_status.as<uint8_t>();
  if (f->opt_timing) {
    while ((int)f->seg_ev.size() < f->n_seg_ev + 2) {
      cudaEvent_t e;
      LH_CUDA(cudaEventCreate(&e));
      f->seg_ev.push_back(e);
    }
    cudaEventRecord(f->seg_ev[f->n_seg_ev], ps);
  }
  lhb::launch_newick(p, ps);
  if (f->opt_timing) {
    cudaEventRecord(f->seg_ev[f->n_seg_ev + 1], ps);
    f->n_seg_ev += 2;
  }
  f->tm.kernel_launches += p.kernel != 1 ? 2 : 1;
  LH_CUDA(cudaEventRecord(f->parsed_ev[ci & 1], ps));
  return 0;
}

// Substitution parameters of the whole batch (10 + R doubles per tree) -> HBM on `s`.
int stage_params(lh_family* f, const double* er, const double* pi, const double* rates, cudaStream_t s) {
  const size_t B = (size_t)f->B, R = (size_t)f->R;
  if (f->misc_busy) LH_CUDA(cudaEventSynchronize(f->misc_ev));
  double* misc = f->h_misc.as<double>();
  stream_copy(misc, er, sizeof(double) * 6 * B);
  stream_copy(misc + 6 * B, pi, sizeof(double) * 4 * B);
  // (rates == nullptr: they are computed and copied chunk by chunk, stage_chunk_rates)
  if (rates) stream_copy(misc + 10 * B, rates, sizeof(double) * R * B);
  const size_t n = rates ? B * (10 + R) : B * 10;
  LH_CUDA(cudaMemcpyAsync(f->d_er.p, misc, n * 8, cudaMemcpyHostToDevice, s));
  LH_CUDA(cudaEventRecord(f->misc_ev, s));
  f->misc_busy = true;
  f->tm.h2d_bytes += (int64_t)(n * 8);
  return 0;
}

// Discrete-Gamma category rates of trees [c0, c0 + cn) from their alpha, on the handle's host threads (the same
// doubles as pll_compute_gamma_cats, host/gamma.cpp; src/PhyloHMM.cpp:425-426), into `rates` and on to the device:
// done chunk by chunk so that it overlaps the kernels of the chunks before instead of standing in front of the call.
int stage_chunk_rates(lh_family* f, const double* alpha, double* rates, int c0, int cn, cudaStream_t s) {
  const size_t B = (size_t)f->B, R = (size_t)f->R;
  ensure_pool(f);
  std::atomic<int> bad(-1);
  std::vector<std::string> msgs((size_t)f->pool->workers() + 1);
  f->pool->run((cn + 63) / 64, [&](int task, int w) {
    const int lo = c0 + task * 64, n = std::min(c0 + cn, lo + 64) - lo;
    if (lhh_gamma_cats_batch(n, alpha + lo, (int)R, rates + (size_t)lo * R)) {
      bad.store(lo);
      msgs[w] = lhh_last_error();
    }
  });
  if (bad.load() >= 0) {
    for (auto& m : msgs)
      if (!m.empty()) return fail("%s (trees %d..)", m.c_str(), bad.load());
    return fail("Gamma category rates failed");
  }
  double* misc = f->h_misc.as<double>() + 10 * B + (size_t)c0 * R;   // (the call's stage_params waited for the buffer)
  stream_copy(misc, rates + (size_t)c0 * R, sizeof(double) * R * cn);
  LH_CUDA(cudaMemcpyAsync(f->d_er.as<double>() + 10 * B + (size_t)c0 * R, misc, sizeof(double) * R * cn, cudaMemcpyHostToDevice, s));
  LH_CUDA(cudaEventRecord(f->misc_ev, s));
  f->tm.h2d_bytes += (int64_t)(sizeof(double) * R * cn);
  return 0;
}

int ensure_chunk_events(lh_family* f, int n_chunks) {
  while ((int)f->chunk_ev.size() < kChunkEvents * n_chunks) {
    cudaEvent_t e;
    LH_CUDA(cudaEventCreate(&e));
    f->chunk_ev.push_back(e);
  }
  while ((int)f->staged_ev.size() < n_chunks) {
    cudaEvent_t e;
    LH_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    f->staged_ev.push_back(e);
  }
  while ((int)f->edge_ev.size() < n_chunks) {
    cudaEvent_t e;
    LH_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    f->edge_ev.push_back(e);
  }
  return 0;
}

// The four kernels of one chunk of trees, in stream order: eigen systems, branch records,
// pruning (xMSA emission), HMM forward pass.
void launch_plan_trees(lh_family* f, int t0, int n, cudaStream_t s) {
  lhb::PlanDev p{};
  p.T = f->fd.T;
  p.tree0 = t0;
  p.children = f->d_rchildren.as<int32_t>();
  p.blen = f->d_rblen.as<double>();
  p.root_child = f->d_rroot_child.as<int32_t>();
  p.root_blen = f->d_rroot_blen.as<double>();
  p.plan = f->d_plan.as<int4>();
  p.pblen = f->d_pblen.as<double>();
  p.flags = f->d_flags.as<int32_t>();
  lhb::launch_plan(p, n, s);
  f->flags_pending = true;
}

// Start values of the k_plan flags -> HBM on `s` (before any k_plan of the batch).
int stage_flags(lh_family* f, cudaStream_t s) {
  if (!f->dev_plan) return 0;
  int32_t* h = f->h_flags.as<int32_t>();
  h[0] = h[1] = INT32_MAX;
  h[2] = h[3] = 0;
  LH_CUDA(cudaMemcpyAsync(f->d_flags.p, h, 4 * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  return 0;
}

void launch_chunk(lh_family* f, int ci, int t0, int n) {
  cudaStream_t s = f->stream;
  const bool tm = f->opt_timing;
  cudaEvent_t* ev = f->chunk_ev.data() + (size_t)kChunkEvents * ci;
  if (tm) cudaEventRecord(ev[0], s);
  lhb::launch_eigen(f->bd, t0, n, s);
  lhb::launch_pmatrix(f->fd, f->bd, t0, n, s);
  if (tm) cudaEventRecord(ev[1], s);
  // trees on site patterns are listed first in the chunk's part of d_list, the others after
  // (text input: the split is made on the device by k_split; both launches go out and each takes its part)
  const bool dyn = f->newick_mode;
  const int n_pat = dyn ? 0 : f->chunk_n_pat[ci];
  const int count[2] = {dyn ? (f->pc[0].n_cols ? n : 0) : n - n_pat, dyn ? (f->pc[1].n_cols ? n : 0) : n_pat};
  const int32_t* list[2] = {f->d_list.as<int32_t>() + t0 + n_pat, f->d_list.as<int32_t>() + t0};
  if (tm) cudaEventRecord(ev[2], s);
  for (int g = 1; g >= 0; --g) {
    if (!count[g]) continue;
    lhb::PruneCfg c = f->pc[g];
    c.tree_list = list[g];
    c.work_counter = f->d_counters.as<int32_t>() + g;
    c.grid = (int)std::min<int64_t>(c.grid, (int64_t)count[g] * c.n_tiles * f->R);
    if (dyn) {
      c.dyn_n_pat = f->d_chunk_npat.as<int32_t>() + ci;
      c.dyn_total = n;
    }
    lhb::launch_prune(f->fd, f->bd, c, t0, count[g], s);
    f->tm.kernel_launches += 1;
    f->tm.prune_launches += 1;
  }
  if (tm) cudaEventRecord(ev[3], s);
  lhb::launch_edge(f->fd, f->bd, f->ec, t0, n, s);
  if (tm) cudaEventRecord(ev[4], s);
  // the forward pass goes to its own stream: the next chunk's kernels do not wait for it
  cudaStream_t hs = f->opt_hmm_overlap ? f->hmm_stream : s;
  if (hs != s) {
    cudaEventRecord(f->edge_ev[ci], s);
    cudaStreamWaitEvent(hs, f->edge_ev[ci], 0);
  }
  if (tm) cudaEventRecord(ev[5], hs);
  lhb::launch_hmm(f->fd, f->bd, f->ho, t0, n, hs);
  if (tm) cudaEventRecord(ev[6], hs);
  f->tm.kernel_launches += 4;
}

// Makes the compute stream wait for the forward passes queued on the HMM stream.
void join_hmm_stream(lh_family* f) {
  if (!f->opt_hmm_overlap) return;
  cudaEventRecord(f->hmm_done, f->hmm_stream);
  cudaStreamWaitEvent(f->stream, f->hmm_done, 0);
}

// Fixes the chunk boundaries of the batch and queues the per-chunk tree lists on `s`.
int stage_lists(lh_family* f, std::vector<int> starts, cudaStream_t s) {
  f->starts = std::move(starts);
  f->n_chunks = (int)f->starts.size() - 1;
  f->chunk_n_pat.assign(f->n_chunks, 0);
  if (f->newick_mode) {  // the lists are made on the device, chunk by chunk (k_split)
    if (f->d_chunk_npat.reserve((size_t)f->n_chunks * sizeof(int32_t)) || f->h_chunk_npat.reserve((size_t)f->n_chunks * sizeof(int32_t)))
      return 1;
    return ensure_chunk_events(f, f->n_chunks);
  }
  if (f->list_busy) LH_CUDA(cudaEventSynchronize(f->list_ev));
  if (f->h_list.reserve((size_t)f->B * sizeof(int32_t))) return 1;
  int32_t* l = f->h_list.as<int32_t>();
  for (int ci = 0; ci < f->n_chunks; ++ci) {
    int w = f->starts[ci];
    for (int t = f->starts[ci]; t < f->starts[ci + 1]; ++t)
      if (f->on_patterns[t]) l[w++] = t;
    f->chunk_n_pat[ci] = w - f->starts[ci];
    for (int t = f->starts[ci]; t < f->starts[ci + 1]; ++t)
      if (!f->on_patterns[t]) l[w++] = t;
  }
  LH_CUDA(cudaMemcpyAsync(f->d_list.p, l, (size_t)f->B * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  LH_CUDA(cudaEventRecord(f->list_ev, s));
  f->list_busy = true;
  f->tm.h2d_bytes += (int64_t)f->B * 4;
  return ensure_chunk_events(f, f->n_chunks);
}

// Chunk boundaries: 16-wave chunks for a resident batch; the pipelined call opens with chunks of
// one and two waves (the first kernels start after a sliver of the host planning) and goes on in
// chunks of `pipe_waves` waves. A batch too small for that (a shard of a multi-GPU job) is cut
// into about four chunks of whole waves, so that planning and kernels still overlap.
std::vector<int> chunk_starts(const lh_family* f, bool pipelined) {
  std::vector<int> st;
  int t0 = 0;
  int step = pipelined ? f->chunk_pipe : f->chunk;
  if (pipelined && f->wave_pipe > 0) {
    // The very first chunk is one item per resident CTA (`wave_unit` trees): the GPU starts after the
    // staging of a few hundred trees; what that short launch loses is less than the lead-in it saves.
    const int w0 = f->opt_open_unit && f->wave_unit > 0 && f->wave_unit * 2 <= f->wave_pipe ? f->wave_unit : 0;
    if (f->chunk_pipe >= 4 * f->wave_pipe && f->B >= 2 * f->chunk_pipe) {
      st.push_back(0);
      if (w0) st.push_back(w0);
      st.push_back(f->wave_pipe);
      t0 = 3 * f->wave_pipe;
    } else {
      // (but no chunk so short - under ~2 ms of kernels, gauged by its bytes of branch records -
      // that the per-chunk launches and stream hand-offs show)
      const int floor_trees = (int)(((size_t)400 << 20) / f->rec_bytes_tree) / f->wave_pipe * f->wave_pipe;
      step = std::min(step, std::max({f->wave_pipe, f->B / 4 / f->wave_pipe * f->wave_pipe, floor_trees}));
      if (w0 && f->B > 2 * w0) {
        st.push_back(0);
        t0 = w0;
      }
    }
  }
  // Later chunks grow (up to the resident batch's 16 waves) as long as the host stages a chunk faster
  // than the GPU scores the one before it: fewer launches and launch tails over the bulk of the batch.
  double growth = 1.0;
  if (pipelined && f->opt_chunk_growth && f->wave_pipe > 0 && f->opt_chunk_trees == 0)
    growth = std::min(2.0, std::max(1.0, 0.6 * f->stage_ratio));
  for (; t0 < f->B; t0 += step) {
    st.push_back(t0);
    if (growth > 1.0 && st.size() > 2) {
      int next = (int)(step * growth) / f->wave_pipe * f->wave_pipe;
      step = std::max(step, std::min(next, f->chunk / f->wave_pipe * f->wave_pipe));
    }
  }
  st.push_back(f->B);
  return st;
}

}  // namespace

extern "C" {

int lh_batch_upload(lh_family_t* f, const lh_tree_batch_t* tb) {
  if (!f || !tb) return fail("null argument");
  if (!f->bound) return fail("no family bound to the handle (the last lh_family_rebind failed)");
  LH_CUDA(cudaSetDevice(f->device));
  if (check_batch(f, tb)) return 1;
  if (configure_batch(f, tb->n_trees, tb->n_rates)) return 1;
  if (f->B == 0) return 0;
  cudaStream_t s = f->stream;
  if (f->opt_timing) cudaEventRecord(f->ev[0], s);
  if (stage_params(f, tb->er, tb->pi, tb->rates, s) || stage_flags(f, s)) return 1;
  if (stage_lists(f, chunk_starts(f, false), s)) return 1;
  for (int t0 = 0, k = 0; t0 < f->B; t0 += f->seg, ++k)
    if (stage_segment(f, tb, t0, std::min(f->seg, f->B - t0), k % lh_family::kRing, s)) return 1;
  // a resident batch is planned once, here; lh_batch_run re-runs the likelihood kernels only
  f->plan_in_call = false;
  if (f->dev_plan) launch_plan_trees(f, 0, f->B, s);
  if (f->opt_timing) cudaEventRecord(f->ev[1], s);
  return 0;
}

int lh_batch_run(lh_family_t* f) {
  if (!f) return fail("null family");
  if (f->B == 0) return 0;
  LH_CUDA(cudaSetDevice(f->device));
  const std::vector<int>& st = f->starts;
  f->tm.kernel_launches = 0;
  f->tm.prune_launches = 0;
  if (f->opt_timing) cudaEventRecord(f->ev[2], f->stream);
  for (int ci = 0; ci < f->n_chunks; ++ci) launch_chunk(f, ci, st[ci], st[ci + 1] - st[ci]);
  join_hmm_stream(f);
  if (f->opt_timing) cudaEventRecord(f->ev[4], f->stream);
  f->run_recorded = true;
  LH_CUDA(cudaGetLastError());
  return 0;
}

// Per-kernel device times of the last run (syncs the stream).
static int collect_run_timings(lh_family* f) {
  if (!f->opt_timing || f->B == 0 || f->n_chunks == 0 || !f->run_recorded) return 0;
  LH_CUDA(cudaEventSynchronize(f->ev[4]));
  // (k_hmm of a chunk overlaps the next chunk's kernels, so the stage times can add up to more than ms_kernels)
  float sum[6] = {}, x = 0;
  static const int from[6] = {0, 1, 2, 3, 5, 7}, to[6] = {1, 2, 3, 4, 6, 8};
  const int n_spans = f->plan_in_call ? 6 : 5;
  f->tm.ms_plan_dev = 0.f;
  for (int ci = 0; ci < f->n_chunks; ++ci)
    for (int k = 0; k < n_spans; ++k) {
      cudaEventElapsedTime(&x, f->chunk_ev[kChunkEvents * ci + from[k]], f->chunk_ev[kChunkEvents * ci + to[k]]);
      sum[k] += x;
    }
  f->tm.ms_plan_dev = sum[5];
  f->tm.ms_newick = 0.f;
  for (int i = 0; i + 1 < f->n_seg_ev; i += 2) {
    cudaEventElapsedTime(&x, f->seg_ev[i], f->seg_ev[i + 1]);
    f->tm.ms_newick += x;
  }
  f->tm.ms_pmatrix = sum[0];
  f->tm.ms_prune = sum[2];
  f->tm.ms_edge = sum[3];
  f->tm.ms_hmm = sum[4];
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
  const bool check_flags = f->dev_plan && f->flags_pending;
  if (check_flags && get(f->h_flags.as<int32_t>() + 4, f->d_flags.p, 4 * sizeof(int32_t))) return 1;
  if (f->newick_mode && (get(f->h_nw_status.p, f->d_nw_status.p, (size_t)B) ||
                         get(f->h_chunk_npat.p, f->d_chunk_npat.p, (size_t)f->n_chunks * sizeof(int32_t))))
    return 1;
  if (o) {
    if (get(o->xmsa_emission, f->d_xe.p, (size_t)B * fd.S * 8)) return 1;
    if (o->site_rate_loglik) {
      if (!f->want_rate_ll) return fail("site_rate_loglik was not produced by the last run (ask for it in lh_score_batch)");
      if (get(o->site_rate_loglik, f->d_rate_ll.p, (size_t)B * f->R * fd.S * 8)) return 1;
    }
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
  if (check_flags) {
    // what the host planner reports while planning, k_plan reports here
    const int32_t* fl = f->h_flags.as<int32_t>() + 4;
    f->flags_pending = false;
    if (fl[0] != INT32_MAX) return fail("tree %d: negative branch length", fl[0]);
    if (fl[1] != INT32_MAX) {
      f->bad_plan_tree = fl[1];
      return fail("tree %d is not a valid post-order operation list", fl[1]);
    }
    f->slots = std::max(f->slots, (int)fl[2]);
    const int slot_bound = std::max(f->pc[0].total_slots, f->pc[1].total_slots);
    if (f->slots > slot_bound) return fail("internal: tree needs %d CLV slots, bound is %d", f->slots, slot_bound);
  }
  std::memcpy(loglik, f->h_out.p, (size_t)B * 8);
  f->tm.d2h_bytes = bytes;
  if (f->newick_mode) {
    f->tm.pattern_trees = 0;
    for (int ci = 0; ci < f->n_chunks; ++ci) f->tm.pattern_trees += f->h_chunk_npat.as<int32_t>()[ci];
  }
  if (f->opt_timing) {
    collect_run_timings(f);
    cudaEventElapsedTime(&f->tm.ms_h2d, f->ev[0], f->ev[1]);
    cudaEventElapsedTime(&f->tm.ms_d2h, f->ev[5], f->ev[6]);
    cudaEventElapsedTime(&f->tm.ms_total, f->ev[0], f->ev[6]);
  }
  return 0;
}

}  // extern "C"

namespace {

// The pipelined call behind lh_score_batch (tb: operation arrays) and lh_score_newick (nb: text; `rates`
// are the category rates of its trees). Host buffers in, host results out.
int score_call(lh_family* f, const lh_tree_batch_t* tb, const lh_newick_batch_t* nb, const double* rates,
               double* loglik, lh_outputs_t* o, double* lazy_rates = nullptr) {  // lazy_rates: [B][R] to fill from nb->alpha
  f->want_tables = f->opt_keep_tables ||
                   (o && (o->vgerm_forward || o->vd_forward || o->dgerm_forward || o->dj_forward ||
                          o->jgerm_forward || o->germ_scalers || o->vd_scalers || o->dj_scalers));
  f->want_rate_ll = o && o->site_rate_loglik;
  LH_CUDA(cudaSetDevice(f->device));
  const auto call_t0 = std::chrono::steady_clock::now();
  auto since = [&]() { return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - call_t0).count(); };
  static const bool trace = getenv("LH_TRACE") != nullptr;
  std::vector<std::pair<const char*, cudaEvent_t>> dev_marks;
  auto mark = [&](const char* what) { if (trace) fprintf(stderr, "LHTRACE %-14s %.3f ms\n", what, since()); };
  auto dmark = [&](const char* what, cudaStream_t st_) {
    if (!trace) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st_);
    dev_marks.emplace_back(what, e);
  };
  f->newick_mode = nb != nullptr;
  f->bad_plan_tree = -1;
  f->n_seg_ev = 0;
  int B, R;
  const double *er, *pi;
  if (nb) {
    B = nb->n_trees;
    R = nb->n_rates;
    er = nb->er;
    pi = nb->pi;
    if (B < 0) return fail("negative batch size");
    if (R < 1 || R > LH_MAX_RATES) return fail("n_rates %d outside 1..%d", R, LH_MAX_RATES);
    for (int t = 0; t < B; ++t)
      for (int i = 0; i < 4; ++i)
        if (!(pi[4 * t + i] > 0.0)) return fail("tree %d: pi[%d] must be positive", t, i);
    // which trees run on site patterns is found by the parser; until then every tree may
    f->on_patterns.clear();
    f->n_pattern_trees = (f->opt_site_patterns && f->cols[1].n_cols < f->cols[0].n_cols) ? B : 0;
  } else {
    if (check_batch(f, tb)) return 1;
    B = tb->n_trees;
    R = tb->n_rates;
    er = tb->er;
    pi = tb->pi;
    rates = tb->rates;
  }
  mark("check");
  if (configure_batch(f, B, R)) return 1;
  if (f->B == 0) return 0;
  f->tm.ms_setup = since();
  mark("configure");
  // Pipeline: the host plans one segment of trees at a time into the pinned ring, the copy
  // stream moves it to HBM, and the compute stream starts a chunk's kernels as soon as the
  // chunk's last segment has landed - planning and H2D of chunk c+1 overlap the kernels of c.
  cudaStream_t cs = f->copy_stream;
  // the copy stream must not overwrite buffers an earlier run on the compute stream still reads
  LH_CUDA(cudaEventRecord(f->ev[7], f->stream));
  LH_CUDA(cudaStreamWaitEvent(cs, f->ev[7], 0));
  if (f->opt_timing) cudaEventRecord(f->ev[0], cs);
  if (stage_params(f, er, pi, rates, cs) || stage_flags(f, cs)) return 1;
  dmark("cs start", cs);
  if (stage_lists(f, chunk_starts(f, true), cs)) return 1;
  mark("params+lists");
  dmark("cs params+lists", cs);
  f->plan_in_call = f->dev_plan;
  // (the plan stream must not overwrite programs an earlier run on the compute stream still reads)
  if (f->dev_plan) LH_CUDA(cudaStreamWaitEvent(f->plan_stream, f->ev[7], 0));
  const std::vector<int>& st = f->starts;
  for (int ci = 0, k = 0; ci < f->n_chunks; ++ci) {
    const int c0 = st[ci], cn = st[ci + 1] - st[ci];
    if (nb) {
      if (lazy_rates && stage_chunk_rates(f, nb->alpha, lazy_rates, c0, cn, cs)) return 1;
      if (stage_text_chunk(f, nb, ci, c0, cn, &k, cs)) return 1;
    } else {
      for (int t0 = c0; t0 < st[ci + 1]; t0 += f->seg, ++k)
        if (stage_segment(f, tb, t0, std::min(f->seg, st[ci + 1] - t0), k % lh_family::kRing, cs)) return 1;
    }
    mark("chunk staged");
    dmark("cs chunk copied", cs);
    LH_CUDA(cudaEventRecord(f->staged_ev[ci], cs));
    if (f->dev_plan) {
      // k_plan of the chunk on its own stream: it runs as soon as the chunk has landed, which is while
      // the chunk before is still in its kernels (in the space k_prune's last CTAs, k_edge and k_hmm leave)
      cudaEvent_t* pev = f->chunk_ev.data() + (size_t)kChunkEvents * ci + 7;
      LH_CUDA(cudaStreamWaitEvent(f->plan_stream, f->staged_ev[ci], 0));
      if (nb) {
        // (the chunk's k_newick launches are already queued on the plan stream)
        lhb::launch_split(f->pc[1].n_cols ? f->d_on_pat.as<uint8_t>() : nullptr, c0, cn, f->d_list.as<int32_t>(),
                          f->d_chunk_npat.as<int32_t>() + ci, f->plan_stream);
        f->tm.kernel_launches += 1;
      }
      LH_CUDA(cudaEventRecord(pev[0], f->plan_stream));
      launch_plan_trees(f, c0, cn, f->plan_stream);
      f->tm.kernel_launches += 1;
      LH_CUDA(cudaEventRecord(pev[1], f->plan_stream));
      LH_CUDA(cudaStreamWaitEvent(f->stream, pev[1], 0));
    } else {
      LH_CUDA(cudaStreamWaitEvent(f->stream, f->staged_ev[ci], 0));
    }
    if (ci == 0 && f->opt_timing) cudaEventRecord(f->ev[2], f->stream);
    dmark("compute chunk start", f->stream);
    launch_chunk(f, ci, c0, cn);
    dmark("compute chunk end", f->stream);
    mark("chunk launched");
  }
  join_hmm_stream(f);
  if (f->opt_timing) {
    cudaEventRecord(f->ev[1], cs);
    cudaEventRecord(f->ev[4], f->stream);
  }
  f->run_recorded = true;
  LH_CUDA(cudaGetLastError());
  const int rc = lh_batch_fetch(f, loglik, o);
  mark("fetched");
  for (size_t i = 0; i < dev_marks.size(); ++i) {
    float x = 0;
    cudaEventElapsedTime(&x, dev_marks[0].second, dev_marks[i].second);
    fprintf(stderr, "LHTRACE dev %-20s %.3f ms\n", dev_marks[i].first, x);
  }
  for (auto& m : dev_marks) cudaEventDestroy(m.second);
  f->tm.ms_call = since();
  return rc;
}

// Copies row `src` of a [n][row] result of the host-parsed trees into row `dst` of the caller's array.
template <class T>
void put_row(T* dst_base, const std::vector<T>& src_base, size_t row, int dst, int src) {
  if (dst_base && row) std::memcpy(dst_base + (size_t)dst * row, src_base.data() + (size_t)src * row, row * sizeof(T));
}

}  // namespace

extern "C" {

int lh_score_batch(lh_family_t* f, const lh_tree_batch_t* tb, double* loglik, lh_outputs_t* o) {
  if (!f || !tb || !loglik) return fail("null argument");
  if (!f->bound) return fail("no family bound to the handle (the last lh_family_rebind failed)");
  return score_call(f, tb, nullptr, nullptr, loglik, o);
}

int lh_score_newick(lh_family_t* f, const lh_newick_batch_t* nb, double* loglik, lh_outputs_t* o) {
  if (!f || !nb || !loglik) return fail("null argument");
  if (!f->bound) return fail("no family bound to the handle (the last lh_family_rebind failed)");
  if (!f->labels_set) return fail("lh_score_newick needs the family's row labels (lh_family_set_labels)");
  const int B = nb->n_trees, R = nb->n_rates, T = f->fd.T;
  if (B > 0 && (!nb->newick || !nb->er || !nb->pi || (!nb->rates && !nb->alpha))) return fail("null argument");
  for (int t = 0; t < B; ++t)
    if (!nb->newick[t]) return fail("tree %d: null string", t);
  // category rates: given, or from alpha on the handle's host threads (the same doubles as
  // pll_compute_gamma_cats, host/gamma.cpp)
  std::vector<double> rates_buf;
  const double* rates = nb->rates;
  if (!rates && B > 0) {
    if (R < 1 || R > LH_MAX_RATES) return fail("n_rates %d outside 1..%d", R, LH_MAX_RATES);
    rates_buf.resize((size_t)B * R);   // filled chunk by chunk inside the pipeline (stage_chunk_rates)
  }
  // text in pinned (or registered) host memory that comes with its lengths can be copied where it lies
  f->text_direct_call = false;
  if (f->opt_text_direct && nb->length && B > 0) {
    LH_CUDA(cudaSetDevice(f->device));
    auto pinned = [](const void* p) {
      cudaPointerAttributes a;
      if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
      }
      return a.type == cudaMemoryTypeHost;
    };
    f->text_direct_call = pinned(nb->newick[0]) && pinned(nb->newick[B - 1] + nb->length[B - 1]);
  }
  int rc = score_call(f, nullptr, nb, rates, loglik, o, rates ? nullptr : rates_buf.data());
  if (!rates) rates = rates_buf.data();
  f->newick_mode = false;
  if (B == 0) return rc;
  const lh_timings_t tm_main = f->tm;
  auto host_parse_message = [&](int t) {
    std::vector<int32_t> ch((size_t)(T - 2) * 2);
    std::vector<double> bl((size_t)(T - 2) * 2);
    int32_t rch;
    double rbl;
    if (lhh_newick_to_ops(T, f->label_ptrs.data(), 1, nb->newick + t, 1, ch.data(), bl.data(), &rch, &rbl)) {
      std::string m = lhh_last_error();
      const std::string pre = "tree 0: ";  // the host parser numbers the trees of its own call
      if (m.compare(0, pre.size(), pre) == 0) m = m.substr(pre.size());
      return m;
    }
    return std::string();
  };
  if (rc) {
    // k_plan refused a tree k_newick had passed (a label twice): the host parser words the error
    if (f->bad_plan_tree >= 0) {
      const std::string m = host_parse_message(f->bad_plan_tree);
      if (!m.empty()) return fail("tree %d: %s", f->bad_plan_tree, m.c_str());
    }
    return rc;
  }
  // trees the device parser left to the host: parse them there (its errors are the call's errors) and
  // score them through the array path; their results replace the stand-ins'
  const uint8_t* status = f->h_nw_status.as<uint8_t>();
  std::vector<int> todo;
  for (int t = 0; t < B; ++t)
    if (status[t]) todo.push_back(t);
  if (todo.empty()) return 0;
  const int n = (int)todo.size();
  const size_t per = (size_t)(T - 2) * 2;
  std::vector<const char*> strs(n);
  for (int i = 0; i < n; ++i) strs[i] = nb->newick[todo[i]];
  std::vector<int32_t> ch(n * per), rch(n);
  std::vector<double> bl(n * per), rbl(n), er(6 * (size_t)n), pi(4 * (size_t)n), rt((size_t)n * R), ll(n);
  if (lhh_newick_to_ops(T, f->label_ptrs.data(), n, strs.data(), f->opt_threads, ch.data(), bl.data(), rch.data(), rbl.data())) {
    // report the first bad tree of the batch with its own message
    for (int i = 0; i < n; ++i) {
      const std::string m = host_parse_message(todo[i]);
      if (!m.empty()) return fail("tree %d: %s", todo[i], m.c_str());
    }
    return fail("%s", lhh_last_error());
  }
  for (int i = 0; i < n; ++i) {
    std::memcpy(&er[6 * (size_t)i], nb->er + 6 * (size_t)todo[i], 6 * sizeof(double));
    std::memcpy(&pi[4 * (size_t)i], nb->pi + 4 * (size_t)todo[i], 4 * sizeof(double));
    std::memcpy(&rt[(size_t)R * i], rates + (size_t)R * todo[i], R * sizeof(double));
  }
  lh_tree_batch_t tb{n, R, ch.data(), bl.data(), rch.data(), rbl.data(), er.data(), pi.data(), rt.data()};
  const lhb::FamilyDev& fd = f->fd;
  const size_t S = fd.S, vd_n = (size_t)fd.vd_sites * fd.vd_states, dj_n = (size_t)fd.dj_sites * fd.dj_states;
  std::vector<double> xe, fv, fvd, fdg, fdj, fj, rll;
  std::vector<int32_t> gs, vds, djs;
  lh_outputs_t o2{};
  if (o) {
    auto want = [&](auto* dst, auto& buf, size_t row) -> decltype(dst) {
      if (!dst || !row) return nullptr;
      buf.resize((size_t)n * row);
      return buf.data();
    };
    o2.xmsa_emission = want(o->xmsa_emission, xe, S);
    o2.vgerm_forward = want(o->vgerm_forward, fv, fd.n_v);
    o2.vd_forward = want(o->vd_forward, fvd, vd_n);
    o2.dgerm_forward = want(o->dgerm_forward, fdg, fd.n_d);
    o2.dj_forward = want(o->dj_forward, fdj, dj_n);
    o2.jgerm_forward = want(o->jgerm_forward, fj, fd.n_j);
    o2.germ_scalers = want(o->germ_scalers, gs, 3);
    o2.vd_scalers = want(o->vd_scalers, vds, fd.vd_sites);
    o2.dj_scalers = want(o->dj_scalers, djs, fd.dj_sites);
    o2.site_rate_loglik = want(o->site_rate_loglik, rll, (size_t)R * S);
  }
  rc = score_call(f, &tb, nullptr, nullptr, ll.data(), o ? &o2 : nullptr);
  if (rc) {
    // (an error of the array path names a tree of the small batch)
    return fail("among the %d trees left to the host parser (first: tree %d): %s", n, todo[0], std::string(g_err).c_str());
  }
  for (int i = 0; i < n; ++i) {
    const int t = todo[i];
    loglik[t] = ll[i];
    if (!o) continue;
    put_row(o->xmsa_emission, xe, o2.xmsa_emission ? S : 0, t, i);
    put_row(o->vgerm_forward, fv, o2.vgerm_forward ? (size_t)fd.n_v : 0, t, i);
    put_row(o->vd_forward, fvd, o2.vd_forward ? vd_n : 0, t, i);
    put_row(o->dgerm_forward, fdg, o2.dgerm_forward ? (size_t)fd.n_d : 0, t, i);
    put_row(o->dj_forward, fdj, o2.dj_forward ? dj_n : 0, t, i);
    put_row(o->jgerm_forward, fj, o2.jgerm_forward ? (size_t)fd.n_j : 0, t, i);
    put_row(o->germ_scalers, gs, o2.germ_scalers ? (size_t)3 : 0, t, i);
    put_row(o->vd_scalers, vds, o2.vd_scalers ? (size_t)fd.vd_sites : 0, t, i);
    put_row(o->dj_scalers, djs, o2.dj_scalers ? (size_t)fd.dj_sites : 0, t, i);
    put_row(o->site_rate_loglik, rll, o2.site_rate_loglik ? (size_t)R * S : 0, t, i);
  }
  // the timings the caller reads are those of the main pass
  const int64_t h2d = f->tm.h2d_bytes, d2h = f->tm.d2h_bytes;
  const int launches = f->tm.kernel_launches;
  f->tm = tm_main;
  f->tm.h2d_bytes += h2d;
  f->tm.d2h_bytes += d2h;
  f->tm.kernel_launches += launches;
  f->tm.newick_host_trees = n;
  f->run_recorded = false;  // (the chunk events now belong to the small batch)
  return 0;
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

// Test hook: the traversal programs of the current batch as the device holds them.
int lh_debug_fetch_plan(lh_family_t* f, int32_t* steps, double* pblen) {
  if (!f || !steps || !pblen) return fail("null argument");
  if (f->B == 0) return 0;
  LH_CUDA(cudaSetDevice(f->device));
  LH_CUDA(cudaStreamSynchronize(f->stream));
  const size_t T = f->fd.T;
  LH_CUDA(cudaMemcpy(steps, f->d_plan.p, (size_t)f->B * (T - 1) * sizeof(lhb::Step), cudaMemcpyDeviceToHost));
  LH_CUDA(cudaMemcpy(pblen, f->d_pblen.p, (size_t)f->B * (2 * T - 3) * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

// Test hook: runs only the device parser over the batch's strings and returns what it wrote.
int lh_debug_parse_newick(lh_family_t* f, const lh_newick_batch_t* nb, int32_t* children, double* blen,
                          int32_t* root_child, double* root_blen, uint8_t* on_patterns, uint8_t* left_to_host, float* ms_kernel) {
  if (!f || !nb || !children || !blen || !root_child || !root_blen || !left_to_host) return fail("null argument");
  if (!f->bound || !f->labels_set) return fail("needs a bound family with labels (lh_family_set_labels)");
  LH_CUDA(cudaSetDevice(f->device));
  const int B = nb->n_trees;
  if (B <= 0) return 0;
  f->newick_mode = true;
  f->n_seg_ev = 0;
  f->on_patterns.clear();
  f->n_pattern_trees = 0;
  const int rc = configure_batch(f, B, 1);
  f->window_set = true;  // (no pruning follows: no probe)
  if (rc) {
    f->newick_mode = false;
    return 1;
  }
  int k = 0;
  const int rc2 = stage_text_chunk(f, nb, 0, 0, B, &k, f->copy_stream);
  f->newick_mode = false;
  if (rc2) return 1;
  LH_CUDA(cudaStreamSynchronize(f->plan_stream));
  for (auto& b : f->ring_busy) b = false;
  const size_t n2 = (size_t)B * (f->fd.T - 2) * 2;
  LH_CUDA(cudaMemcpy(children, f->d_rchildren.p, n2 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  LH_CUDA(cudaMemcpy(blen, f->d_rblen.p, n2 * sizeof(double), cudaMemcpyDeviceToHost));
  LH_CUDA(cudaMemcpy(root_child, f->d_rroot_child.p, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost));
  LH_CUDA(cudaMemcpy(root_blen, f->d_rroot_blen.p, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost));
  if (on_patterns) LH_CUDA(cudaMemcpy(on_patterns, f->d_on_pat.p, (size_t)B, cudaMemcpyDeviceToHost));
  LH_CUDA(cudaMemcpy(left_to_host, f->d_nw_status.p, (size_t)B, cudaMemcpyDeviceToHost));
  if (ms_kernel) {
    *ms_kernel = 0.f;
    for (int i = 0; i + 1 < f->n_seg_ev; i += 2) {
      float x = 0.f;
      cudaEventElapsedTime(&x, f->seg_ev[i], f->seg_ev[i + 1]);
      *ms_kernel += x;
    }
  }
  f->B = 0;  // no batch is resident
  return 0;
}

// The device Newick parser on caller-owned device buffers, with its scratch in a static device array: calls no
// cudaMalloc / cudaFree (nor anything else that allocates) and only queues two kernels on `stream`.
int lh_newick_parse_noalloc(lh_family_t* f, int32_t n_trees, const char* text_dev, const uint64_t* offs_dev,
                            const uint32_t* lens_dev, int32_t* children_dev, double* blen_dev, int32_t* root_child_dev,
                            double* root_blen_dev, uint8_t* on_patterns_dev, uint8_t* status_dev, void* stream) {
  if (!f) return fail("null family");
  if (!f->bound || !f->labels_set) return fail("lh_newick_parse_noalloc needs a bound family with labels (lh_family_set_labels)");
  if (n_trees < 0) return fail("negative tree count");
  const int max_trees = lhb::newick_noalloc_max_trees(f->fd.T);
  if (max_trees == 0) return fail("lh_newick_parse_noalloc: a family of %d tips is beyond the static scratch", f->fd.T);
  if (n_trees > max_trees)
    return fail("lh_newick_parse_noalloc: %d trees of %d tips need more than the %zu-byte static scratch (at most %d trees per call)",
                n_trees, f->fd.T, (size_t)lhb::kNoAllocScratchBytes, max_trees);
  if (n_trees == 0) return 0;
  if (!text_dev || !offs_dev || !lens_dev || !children_dev || !blen_dev || !root_child_dev || !root_blen_dev ||
      !on_patterns_dev || !status_dev)
    return fail("null argument");
  LH_CUDA(cudaSetDevice(f->device));
  lhb::NewickDev p{};
  p.T = f->fd.T;
  p.tree0 = 0;
  p.n = n_trees;
  p.text = text_dev;
  p.offs = offs_dev;
  p.lens = lens_dev;
  p.kernel = 2;
  p.labels = f->lt;
  p.children = children_dev;
  p.blen = blen_dev;
  p.root_child = root_child_dev;
  p.root_blen = root_blen_dev;
  p.on_pat = on_patterns_dev;
  p.status = status_dev;
  const int rc = lhb::launch_newick_noalloc(p, f->device, static_cast<cudaStream_t>(stream));
  if (rc) return fail("lh_newick_parse_noalloc: %s", cudaGetErrorString(static_cast<cudaError_t>(rc)));
  return 0;
}

int32_t lh_newick_parse_noalloc_max_trees(int32_t n_tips) { return lhb::newick_noalloc_max_trees(n_tips); }

// Ancestral-state sampling (asr.cu): the family handle names the device and the tip count; the alignment is the
// caller's (the MSA with the naive row replaced per sample, not the xMSA).
int lh_asr_sample(lh_family_t* f, const lh_asr_batch_t* a, uint8_t* node_states, uint8_t* site_rate,
                  double* site_rate_loglik) {
  if (!f || !a || !a->trees) return fail("null argument");
  if (!f->bound) return fail("lh_asr_sample needs a bound family");
  const lh_tree_batch_t* tb = a->trees;
  const int T = f->fd.T, B = tb->n_trees, R = tb->n_rates, S = a->n_sites;
  if (B < 0 || S < 0) return fail("negative tree or site count");
  if (R < 1 || R > LH_MAX_RATES) return fail("n_rates %d outside 1..%d", R, LH_MAX_RATES);
  if (B == 0 || S == 0) return 0;
  if (!tb->children || !tb->blen || !tb->root_child || !tb->root_blen || !tb->er || !tb->pi || !tb->rates ||
      !a->msa_codes || !a->naive_codes || !node_states || !site_rate)
    return fail("null argument");
  if (a->naive_row < -1 || a->naive_row >= T) return fail("naive_row %d outside -1..%d", a->naive_row, T - 1);
  const int n_ops = T - 2;
  for (int b = 0; b < B; ++b) {
    // children before parents, every inner node used once: what the kernel's single forward loop relies on
    const int32_t* ch = tb->children + (size_t)b * n_ops * 2;
    for (int op = 0; op < n_ops; ++op)
      for (int c = 0; c < 2; ++c) {
        const int id = ch[2 * op + c];
        if (id < 0 || id >= T + op) return fail("tree %d: operation %d refers to node %d before it is computed", b, op, id);
      }
    const int rc = tb->root_child[b];
    if (rc < 0 || rc >= T + n_ops - 1) return fail("tree %d: root_child %d out of range", b, rc);
    for (int i = 0; i < 2 * n_ops; ++i)
      if (!(tb->blen[(size_t)b * n_ops * 2 + i] >= 0.0)) return fail("tree %d: negative branch length", b);
    if (!(tb->root_blen[b] >= 0.0)) return fail("tree %d: negative branch length", b);
  }
  LH_CUDA(cudaSetDevice(f->device));
  std::string err;
  const size_t scratch = f->opt_asr_scratch_mb > 0 ? (size_t)f->opt_asr_scratch_mb << 20 : (size_t)4 << 30;
  if (lhb::asr_sample(&f->asr_cache, T, B, R, tb->children, tb->blen, tb->root_child, tb->root_blen, tb->er, tb->pi, tb->rates, S,
                      a->msa_codes, a->naive_row, a->naive_codes, a->seed, scratch, node_states, site_rate,
                      site_rate_loglik, &err))
    return fail("%s", err.c_str());
  return 0;
}

// The index draws of SampleNaiveSequence for every tree of the batch scored last (k_sample), from the forward tables
// that run left on the device (option "keep_tables", or read-outs requested from it).
int lh_sample_paths(lh_family_t* f, int32_t n_trees, const uint32_t* words, int32_t n_words, int32_t* idx, int32_t n_idx,
                    uint8_t* flags) {
  if (!f || !words || !idx || !flags) return fail("null argument");
  if (!f->bound) return fail("no family bound to the handle");
  if (!f->want_tables || f->B != n_trees)
    return fail("lh_sample_paths: the last run of the handle did not keep the forward tables of %d trees (option keep_tables)", n_trees);
  const lhb::FamilyDev& fd = f->fd;
  auto draw_words = [](int n_weights, int calls) { return n_weights >= 2 ? 2 * calls : 0; };
  int want_words = draw_words(fd.n_j, 1) + draw_words(fd.vd_states, fd.vd_sites) + draw_words(fd.n_v, 1);
  int want_idx = 2 + fd.vd_sites;
  if (fd.has_d) {
    want_words += draw_words(fd.dj_states, fd.dj_sites) + draw_words(fd.n_d, 1);
    want_idx += 1 + fd.dj_sites;
  }
  if (n_words != want_words || n_idx != want_idx)
    return fail("lh_sample_paths: this family draws %d generator outputs and %d indices per tree (got %d, %d)", want_words,
                want_idx, n_words, n_idx);
  if (n_trees == 0) return 0;
  LH_CUDA(cudaSetDevice(f->device));
  const size_t B = (size_t)n_trees;
  if (f->d_words.reserve(B * std::max(n_words, 1) * 4) || f->d_sidx.reserve(B * n_idx * 4) || f->d_sflag.reserve(B)) return 1;
  cudaStream_t s = f->stream;
  if (n_words) LH_CUDA(cudaMemcpyAsync(f->d_words.p, words, B * n_words * 4, cudaMemcpyHostToDevice, s));
  lhb::launch_sample(fd, f->d_vgerm.as<double>(), f->d_vd.as<double>(), f->d_dgerm.as<double>(), f->d_dj.as<double>(),
                     f->d_jgerm.as<double>(), f->d_words.as<uint32_t>(), n_words, f->d_sidx.as<int32_t>(), n_idx,
                     f->d_sflag.as<uint8_t>(), n_trees, s);
  LH_CUDA(cudaGetLastError());
  LH_CUDA(cudaMemcpyAsync(idx, f->d_sidx.p, B * n_idx * 4, cudaMemcpyDeviceToHost, s));
  LH_CUDA(cudaMemcpyAsync(flags, f->d_sflag.p, B, cudaMemcpyDeviceToHost, s));
  LH_CUDA(cudaStreamSynchronize(s));
  return 0;
}

// Page-lock memory the caller already holds (cudaHostRegister, portable: every device of the process may copy from
// it): text in it is copied in place by lh_score_newick. Costs ~0.1 ms per MB once; undo with lh_unpin_host.
int lh_pin_host(void* ptr, size_t bytes) {
  if (!ptr || bytes == 0) return fail("lh_pin_host: null or empty range");
  LH_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
  return 0;
}
int lh_unpin_host(void* ptr) {
  if (!ptr) return fail("lh_unpin_host: null pointer");
  LH_CUDA(cudaHostUnregister(ptr));
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
