// Ancestral-state sampling: the second half of the reference's R step (scripts/run_bootstrap_asr_ess.R:81-89).
// Per kept tree sample and alignment site the script (a) evaluates the site's likelihood on the tree scaled by
// each discrete-Gamma rate with phylomd::phylo.likelihood and samples the site's rate category in proportion,
// then (b) calls phylomd::asr.sim on the tree scaled by that rate: a post-order pass that keeps every node's
// partial likelihoods, the root state drawn from pi[s] * L_root[s], and every other inner node top-down from
// P(parent state -> s) * L_node[s] (lib/phylomd/src/simulation_functions.cpp:57-117). The alignment is the
// family's MSA with the "naive" row replaced by the sample's NaiveSequence (R:66-67).
//
// k_prune keeps no inner CLV (that is its point), so this is a separate, simpler path for the much smaller job
// (the script runs on a bootstrap subsample of the trace): one thread per (tree, site),
//   k_asr_pmat  P(branch, category) = I + V^-1 diag(expm1(lambda r t)) V for the 2T-3 branches of every tree
//               (the arithmetic of k_pmatrix, core_pmatrix.c:188-237), 128 bytes each, row i = parent state
//   k_asr       per thread: R upward passes over the tree's operations (children before parents, as the batch
//               arrays hold them) with the node CLVs in a global scratch [node][state][site] - coalesced over the
//               sites of a tree -, log L_k per category, the category draw, the upward pass of the drawn category
//               once more if it was not the last one, then the downward pass over the same scratch.
// The model is reversible, so the joint distribution of the inner states does not depend on where the tree is
// rooted: the script roots at the naive outgroup, this kernel at the trifurcation the batch arrays end with.
// Rescaling: a node whose four entries are all below 2^-256 is multiplied by 2^256 and counted (as libpll does
// per site, core_partials.c:648-659); only log L_k needs the count, the conditionals of the downward pass are
// ratios within one node.
// Random numbers: Philox4x32-10 keyed by the caller's seed, counter = (tree index in the call, site, draw number):
// draw 0 is the rate category, draw 1 + n the state of inner node n. Results therefore do not depend on how the
// call is cut into chunks or which GPU runs it, and the CPU restatement the tests hold reproduces them exactly. The reference
// draws with R's generator on a cluster of workers (R:43-46), which nothing outside R reproduces: parity with the
// reference is distributional (tests/test_gpu_asr.py checks the draws against the exact posterior of a small
// tree), parity with the oracle is exact.
// Every product and sum below is written with __dmul_rn / __dadd_rn in a fixed order so that no FMA contraction
// makes the device differ from the oracle's plain IEEE arithmetic.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

#include "kernels.hpp"
#include "philox.hpp"

namespace lhb {
namespace {

constexpr int kAsrMaxRates = 8;   // LH_MAX_RATES (include/linearham_b200.h)

struct AsrDev {
  int32_t T, S, R, n;          // tips, alignment sites, rate categories, trees of this chunk
  int32_t tree0;               // index of the chunk's first tree in the call (RNG counters)
  const int32_t* children;     // [n][T-2][2]
  const double* blen;          // [n][T-2][2]
  const int32_t* root_child;   // [n]
  const double* root_blen;     // [n]
  const double* eig;           // [n][kEigStride]
  const double* rates;         // [n][R]
  const uint8_t* msa;          // [T][S]
  const uint8_t* naive;        // [n][S]
  int32_t naive_row;
  double* pm;                  // [n][2T-3][R][16]
  double* clv;                 // [n][T-2][4][S]
  uint8_t* states;             // [n][T-2][S]
  uint8_t* site_rate;          // [n][S]
  double* srl;                 // [n][R][S] or nullptr
  uint32_t seed_lo, seed_hi;
};

__global__ void __launch_bounds__(128) k_asr_pmat(AsrDev a) {
  const int NB = 2 * a.T - 3;
  const long long total = (long long)a.n * NB * a.R;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(g % a.R);
    const int br = (int)((g / a.R) % NB);
    const int lt = (int)(g / ((long long)a.R * NB));
    const double* e = a.eig + (size_t)lt * kEigStride;
    const double len = br < NB - 1 ? a.blen[(size_t)lt * (NB - 1) + br] : a.root_blen[lt];
    const double rate = a.rates[(size_t)lt * a.R + k];
    double* P = a.pm + (size_t)g * 16;
    if (len == 0.0) {
      for (int j = 0; j < 4; ++j)
        for (int c = 0; c < 4; ++c) P[j * 4 + c] = (j == c) ? 1.0 : 0.0;
      continue;
    }
    double expd[4];
    for (int m = 0; m < 4; ++m) expd[m] = expm1(e[m] * rate * len);
    for (int j = 0; j < 4; ++j) {
      double tmp[4];
      for (int m = 0; m < 4; ++m) tmp[m] = e[20 + j * 4 + m] * expd[m];
      for (int c = 0; c < 4; ++c) {
        double acc = (j == c) ? 1.0 : 0.0;
        for (int m = 0; m < 4; ++m) acc += tmp[m] * e[4 + m * 4 + c];
        P[j * 4 + c] = acc;
      }
    }
  }
}

__device__ __forceinline__ double sum4(double a, double b, double c, double d) {
  return __dadd_rn(__dadd_rn(__dadd_rn(a, b), c), d);
}

// message of a child through its branch: m[i] = sum_j P[i][j] * v[j]; a tip's v is one-hot, or all ones for N / gap
__device__ __forceinline__ void message(const double* __restrict__ P, const double (&v)[4], double (&m)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    m[i] = sum4(__dmul_rn(P[i * 4 + 0], v[0]), __dmul_rn(P[i * 4 + 1], v[1]), __dmul_rn(P[i * 4 + 2], v[2]),
                __dmul_rn(P[i * 4 + 3], v[3]));
}
__device__ __forceinline__ void tip_message(const double* __restrict__ P, int code, double (&m)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    m[i] = code < 4 ? P[i * 4 + code] : sum4(P[i * 4 + 0], P[i * 4 + 1], P[i * 4 + 2], P[i * 4 + 3]);
}

// first index whose running sum exceeds u * total; 4 when the weights do not sum to a positive finite number
__device__ __forceinline__ int draw4(const double (&w)[4], double u) {
  const double c1 = __dadd_rn(w[0], w[1]), c2 = __dadd_rn(c1, w[2]), total = __dadd_rn(c2, w[3]);
  if (!(total > 0.0) || !(total < INFINITY)) return 4;
  const double thr = __dmul_rn(u, total);
  return thr < w[0] ? 0 : thr < c1 ? 1 : thr < c2 ? 2 : 3;
}

__global__ void __launch_bounds__(128) k_asr(AsrDev a) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, lt = blockIdx.y;
  if (j >= a.S) return;
  const int T = a.T, S = a.S, R = a.R, n_ops = T - 2, NB = 2 * T - 3;
  const int32_t* ch = a.children + (size_t)lt * n_ops * 2;
  const double* pm = a.pm + (size_t)lt * NB * R * 16;
  double* clv = a.clv + (size_t)lt * n_ops * 4 * S + j;   // [op][state] at stride S
  uint8_t* st = a.states + (size_t)lt * n_ops * S + j;     // [op] at stride S
  const double* e = a.eig + (size_t)lt * kEigStride;
  const double pi[4] = {e[40], e[41], e[42], e[43]};
  const int root_child = a.root_child[lt];
  auto code_of = [&](int tip) -> int {
    return tip == a.naive_row ? (int)a.naive[(size_t)lt * S + j] : (int)a.msa[(size_t)tip * S + j];
  };
  auto child_message = [&](int id, const double* P, double (&m)[4]) {
    if (id < T) {
      tip_message(P, code_of(id), m);
    } else {
      const double* c = clv + (size_t)(id - T) * 4 * S;
      const double v[4] = {c[0], c[S], c[2 * (size_t)S], c[3 * (size_t)S]};
      message(P, v, m);
    }
  };
  // one upward pass for category k; returns log L_k of the site
  auto upward = [&](int k) -> double {
    int count = 0;
    double node[4];
    for (int op = 0; op < n_ops; ++op) {
      double m0[4], m1[4];
      child_message(ch[2 * op], pm + ((size_t)(2 * op) * R + k) * 16, m0);
      child_message(ch[2 * op + 1], pm + ((size_t)(2 * op + 1) * R + k) * 16, m1);
#pragma unroll
      for (int i = 0; i < 4; ++i) node[i] = __dmul_rn(m0[i], m1[i]);
      if (op == n_ops - 1) {   // the trifurcation: the third neighbour across the root edge
        double m2[4];
        child_message(root_child, pm + ((size_t)(NB - 1) * R + k) * 16, m2);
#pragma unroll
        for (int i = 0; i < 4; ++i) node[i] = __dmul_rn(node[i], m2[i]);
      }
      double mx = fmax(fmax(node[0], node[1]), fmax(node[2], node[3]));
      while (mx < 0x1p-256 && mx > 0.0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) node[i] = __dmul_rn(node[i], 0x1p+256);
        mx = __dmul_rn(mx, 0x1p+256);
        ++count;
      }
      double* c = clv + (size_t)op * 4 * S;
      c[0] = node[0];
      c[S] = node[1];
      c[2 * (size_t)S] = node[2];
      c[3 * (size_t)S] = node[3];
    }
    const double L = sum4(__dmul_rn(pi[0], node[0]), __dmul_rn(pi[1], node[1]), __dmul_rn(pi[2], node[2]),
                          __dmul_rn(pi[3], node[3]));
    return log(L) - (double)count * 177.445678223345993274;   // 256 ln 2
  };

  // (a) the rate category of the site, in proportion to L_k (the division by pi[naive base] of R:85 is common to
  // the categories; it only shows in the read-out)
  double logl[kAsrMaxRates];
  double best = -INFINITY;
  for (int k = 0; k < R; ++k) {
    logl[k] = upward(k);
    best = fmax(best, logl[k]);
  }
  const int naive_code = (int)a.naive[(size_t)lt * S + j];
  if (a.srl) {
    const double lp = naive_code < 4 ? e[36 + naive_code] : 0.0;
    for (int k = 0; k < R; ++k) a.srl[((size_t)lt * R + k) * S + j] = logl[k] - lp;
  }
  const uint32_t tree = (uint32_t)(a.tree0 + lt);
  int k_samp = R;   // R: no category has a positive likelihood
  if (best > -INFINITY && best < INFINITY) {
    double total = 0.0;
    for (int k = 0; k < R; ++k) total = __dadd_rn(total, exp(logl[k] - best));
    const double thr = __dmul_rn(philox_uniform(philox4x32_10(tree, (uint32_t)j, 0u, 0u, a.seed_lo, a.seed_hi).v[0]), total);
    double run = 0.0;
    k_samp = R - 1;
    for (int k = 0; k < R; ++k) {
      run = __dadd_rn(run, exp(logl[k] - best));
      if (thr < run) {
        k_samp = k;
        break;
      }
    }
  }
  a.site_rate[(size_t)lt * S + j] = (uint8_t)k_samp;
  if (k_samp == R) {
    for (int op = 0; op < n_ops; ++op) st[(size_t)op * S] = 4;
    return;
  }
  if (k_samp != R - 1) upward(k_samp);   // the scratch holds the last pass

  // (b) downward: the trifurcation from pi[s] * L[s], every other inner node from P(parent -> s) * L_node[s]
  {
    const double* c = clv + (size_t)(n_ops - 1) * 4 * S;
    const double w[4] = {__dmul_rn(pi[0], c[0]), __dmul_rn(pi[1], c[S]), __dmul_rn(pi[2], c[2 * (size_t)S]),
                         __dmul_rn(pi[3], c[3 * (size_t)S])};
    const double u = philox_uniform(philox4x32_10(tree, (uint32_t)j, (uint32_t)n_ops, 0u, a.seed_lo, a.seed_hi).v[0]);
    st[(size_t)(n_ops - 1) * S] = (uint8_t)draw4(w, u);
  }
  auto draw_child = [&](int id, const double* P, int parent_state) {
    if (id < T) return;
    const int node = id - T;
    int s = 4;
    if (parent_state < 4) {
      const double* c = clv + (size_t)node * 4 * S;
      const double w[4] = {__dmul_rn(P[parent_state * 4 + 0], c[0]), __dmul_rn(P[parent_state * 4 + 1], c[S]),
                           __dmul_rn(P[parent_state * 4 + 2], c[2 * (size_t)S]),
                           __dmul_rn(P[parent_state * 4 + 3], c[3 * (size_t)S])};
      const double u = philox_uniform(philox4x32_10(tree, (uint32_t)j, (uint32_t)(1 + node), 0u, a.seed_lo, a.seed_hi).v[0]);
      s = draw4(w, u);
    }
    st[(size_t)node * S] = (uint8_t)s;
  };
  for (int op = n_ops - 1; op >= 0; --op) {
    const int ps = st[(size_t)op * S];
    if (op == n_ops - 1) draw_child(root_child, pm + ((size_t)(NB - 1) * R + k_samp) * 16, ps);
    draw_child(ch[2 * op], pm + ((size_t)(2 * op) * R + k_samp) * 16, ps);
    draw_child(ch[2 * op + 1], pm + ((size_t)(2 * op + 1) * R + k_samp) * 16, ps);
  }
}

// One device allocation per handle, kept between calls and carved into the call's buffers (a call used to spend
// more time in cudaMalloc / cudaFree of its 3 GB scratch than in its kernels).
struct AsrPool {
  unsigned char* base = nullptr;
  size_t cap = 0, at = 0;
  template <class T>
  T* take(size_t n) {
    T* p = reinterpret_cast<T*>(base + at);
    at += (std::max<size_t>(n, 1) * sizeof(T) + 255) / 256 * 256;
    return p;
  }
};

}  // namespace

#define LH_ASR_CK(call)                                                                 \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      *err = std::string("lh_asr_sample: ") + #call + ": " + cudaGetErrorString(e_);    \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

void asr_cache_free(void* cache) {
  AsrPool* pool = static_cast<AsrPool*>(cache);
  if (!pool) return;
  if (pool->base) cudaFree(pool->base);
  delete pool;
}

// Host driver: chunks of trees sized so that the CLV scratch stays below `scratch_bytes`; synchronous. *cache holds
// the device pool between calls (owned by the family handle, released with asr_cache_free).
int asr_sample(void** cache, int n_tips, int n_trees, int n_rates, const int32_t* children, const double* blen,
               const int32_t* root_child, const double* root_blen, const double* er, const double* pi,
               const double* rates, int n_sites, const uint8_t* msa, int naive_row, const uint8_t* naive,
               uint64_t seed, size_t scratch_bytes, uint8_t* node_states, uint8_t* site_rate, double* site_rate_loglik,
               std::string* err) {
  const int T = n_tips, S = n_sites, R = n_rates, n_ops = T - 2, NB = 2 * T - 3;
  const size_t per_tree = (size_t)n_ops * 4 * S * sizeof(double);
  const int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)n_trees, scratch_bytes / std::max<size_t>(per_tree, 1)));
  if (!*cache) *cache = new AsrPool();
  AsrPool& pool = *static_cast<AsrPool*>(*cache);
  struct { int32_t *p; } d_children, d_root_child;
  struct { double* p; } d_blen, d_root_blen, d_er, d_pi, d_rates, d_eig, d_pm, d_clv, d_srl;
  struct { uint8_t* p; } d_msa, d_naive, d_states, d_site_rate;
  for (int pass = 0; pass < 2; ++pass) {   // pass 0 measures, pass 1 carves
    pool.at = 0;
    d_children.p = pool.take<int32_t>((size_t)chunk * n_ops * 2);
    d_blen.p = pool.take<double>((size_t)chunk * n_ops * 2);
    d_root_child.p = pool.take<int32_t>(chunk);
    d_root_blen.p = pool.take<double>(chunk);
    d_er.p = pool.take<double>((size_t)chunk * 6);
    d_pi.p = pool.take<double>((size_t)chunk * 4);
    d_rates.p = pool.take<double>((size_t)chunk * R);
    d_eig.p = pool.take<double>((size_t)chunk * kEigStride);
    d_pm.p = pool.take<double>((size_t)chunk * NB * R * 16);
    d_clv.p = pool.take<double>((size_t)chunk * n_ops * 4 * S);
    d_msa.p = pool.take<uint8_t>((size_t)T * S);
    d_naive.p = pool.take<uint8_t>((size_t)chunk * S);
    d_states.p = pool.take<uint8_t>((size_t)chunk * n_ops * S);
    d_site_rate.p = pool.take<uint8_t>((size_t)chunk * S);
    d_srl.p = site_rate_loglik ? pool.take<double>((size_t)chunk * R * S) : nullptr;
    if (pass == 0 && pool.at > pool.cap) {
      if (pool.base) LH_ASR_CK(cudaFree(pool.base));
      pool.base = nullptr;
      pool.cap = 0;
      LH_ASR_CK(cudaMalloc(&pool.base, pool.at));
      pool.cap = pool.at;
    }
  }
  LH_ASR_CK(cudaMemcpy(d_msa.p, msa, (size_t)T * S, cudaMemcpyHostToDevice));
  for (int t0 = 0; t0 < n_trees; t0 += chunk) {
    const int n = std::min(chunk, n_trees - t0);
    LH_ASR_CK(cudaMemcpy(d_children.p, children + (size_t)t0 * n_ops * 2, (size_t)n * n_ops * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_blen.p, blen + (size_t)t0 * n_ops * 2, (size_t)n * n_ops * 2 * sizeof(double), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_root_child.p, root_child + t0, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_root_blen.p, root_blen + t0, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_er.p, er + (size_t)t0 * 6, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_pi.p, pi + (size_t)t0 * 4, (size_t)n * 4 * sizeof(double), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_rates.p, rates + (size_t)t0 * R, (size_t)n * R * sizeof(double), cudaMemcpyHostToDevice));
    LH_ASR_CK(cudaMemcpy(d_naive.p, naive + (size_t)t0 * S, (size_t)n * S, cudaMemcpyHostToDevice));
    BatchDev b{};
    b.B = n;
    b.R = R;
    b.er = d_er.p;
    b.pi = d_pi.p;
    b.rates = d_rates.p;
    b.eig = d_eig.p;
    launch_eigen(b, 0, n, 0);
    AsrDev a{};
    a.T = T; a.S = S; a.R = R; a.n = n; a.tree0 = t0;
    a.children = d_children.p; a.blen = d_blen.p; a.root_child = d_root_child.p; a.root_blen = d_root_blen.p;
    a.eig = d_eig.p; a.rates = d_rates.p; a.msa = d_msa.p; a.naive = d_naive.p; a.naive_row = naive_row;
    a.pm = d_pm.p; a.clv = d_clv.p; a.states = d_states.p; a.site_rate = d_site_rate.p;
    a.srl = site_rate_loglik ? d_srl.p : nullptr;
    a.seed_lo = (uint32_t)seed;
    a.seed_hi = (uint32_t)(seed >> 32);
    const long long pm_threads = (long long)n * NB * R;
    k_asr_pmat<<<(unsigned)std::min<long long>((pm_threads + 127) / 128, 148LL * 32), 128>>>(a);
    k_asr<<<dim3((unsigned)((S + 127) / 128), (unsigned)n), 128>>>(a);
    LH_ASR_CK(cudaGetLastError());
    LH_ASR_CK(cudaMemcpy(node_states + (size_t)t0 * n_ops * S, d_states.p, (size_t)n * n_ops * S, cudaMemcpyDeviceToHost));
    LH_ASR_CK(cudaMemcpy(site_rate + (size_t)t0 * S, d_site_rate.p, (size_t)n * S, cudaMemcpyDeviceToHost));
    if (site_rate_loglik)
      LH_ASR_CK(cudaMemcpy(site_rate_loglik + (size_t)t0 * R * S, d_srl.p, (size_t)n * R * S * sizeof(double), cudaMemcpyDeviceToHost));
  }
  return 0;
}

}  // namespace lhb
