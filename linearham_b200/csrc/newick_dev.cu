// k_newick: the Newick parser on the device. Replaces, per tree sample,
//   std::regex_replace("[&index=N]") + pll_utree_parse_newick_string   src/PhyloHMM.cpp:419-421
//   set_missing_branch_length(tree, EPS)                               pll_util.cpp:46-62
//   pll_utree_traverse(POSTORDER) + pll_utree_create_operations         utree.c:284-329, 379-442
// with the same conventions as the host parser (host/newick.cpp, which stays the authority: every
// tree this kernel does not accept - malformed, an unusual number spelling, more than 19 significant
// digits - is flagged and parsed there, so accept/reject decisions and error messages are the host's).
//
// One thread per tree; the trees of a batch are the parallelism (a 1001-tip string is 73 KB and the
// parse is a dependent chain over its characters). A thread reads its string eight bytes at a time into a
// register window. There is no node graph and no scratch memory: the stack of open parentheses lives in
// the tree's own output rows - operations are emitted front to back, and a node that is still open will
// become an operation behind every operation emitted so far, so level d of the stack sits in row
// n_ops - d and is consumed before an emitted operation can reach it.
// Branch lengths are converted with the Eisel-Lemire algorithm (decimal.hpp): correctly rounded, hence
// the same doubles as atof / from_chars.
#include <cuda_runtime.h>

#include <cstdint>

#include "decimal.hpp"
#include "kernels.hpp"

namespace lhb {
namespace {

constexpr double kEpsLen = 1e-6;  // src/utils.hpp:20

struct Reader {
  const char* s;
  uint64_t cur;
  uint32_t pos;
  __device__ __forceinline__ explicit Reader(const char* str) : s(str), pos(0) {
    cur = __ldg(reinterpret_cast<const unsigned long long*>(s));
  }
  __device__ __forceinline__ int peek() const { return (int)((cur >> ((pos & 7u) * 8u)) & 0xffu); }
  __device__ __forceinline__ void adv() {
    ++pos;
    if ((pos & 7u) == 0u) cur = __ldg(reinterpret_cast<const unsigned long long*>(s + pos));
  }
};

// ( ) , : ; [ blank tab newline cr nul
__device__ __forceinline__ bool is_delim(int c) {
  constexpr uint64_t m = (1ull << '(') | (1ull << ')') | (1ull << ',') | (1ull << ':') | (1ull << ';') | (1ull << ' ') |
                         (1ull << '\t') | (1ull << '\n') | (1ull << '\r') | 1ull;
  return (c < 64 && ((m >> c) & 1ull)) || c == '[';
}

// blanks and [...] comments (the reference strips "[&index=N]" with a regex)
__device__ __forceinline__ void skip(Reader& r) {
  for (;;) {
    int c = r.peek();
    while (c == ' ' || c == '\t' || c == '\n' || c == '\r') {
      r.adv();
      c = r.peek();
    }
    if (c != '[') return;
    while (c != ']' && c != 0) {
      r.adv();
      c = r.peek();
    }
    if (c != 0) r.adv();
  }
}

// optional ":length"; a zero or missing length becomes EPS. false: leave the tree to the host
__device__ __forceinline__ bool branch_length(Reader& r, double* len) {
  skip(r);
  double v = 0.0;
  if (r.peek() == ':') {
    r.adv();
    skip(r);
    const int st = parse_decimal([&]() { return r.peek(); }, [&]() { r.adv(); }, &v);
    if (st != kDecimalOk || !is_delim(r.peek())) return false;
  }
  *len = v != 0.0 ? v : kEpsLen;
  return true;
}

__device__ int find_label(const LabelTableDev& lt, const char* text, uint32_t start, int len, uint32_t h, uint64_t first8) {
  const uint32_t lo = (uint32_t)first8, hi = (uint32_t)(first8 >> 32);
  for (uint32_t j = h & lt.mask;; j = (j + 1) & lt.mask) {
    const uint4 e = __ldg(lt.slots + j);
    if (e.w == 0xffffffffu) return -1;
    if (e.z != h || (int)(e.w & 0xffffu) != len || e.x != lo || e.y != hi) continue;
    const int row = (int)(e.w >> 16);
    if (len > 8) {
      const char* ref = lt.chars + __ldg(lt.long_off + row);
      bool same = true;
      for (int i = 8; i < len && same; ++i) same = text[start + i] == ref[i];
      if (!same) continue;
    }
    return row;
  }
}

__global__ void __launch_bounds__(32) k_newick(NewickDev p) {
  const int i = blockIdx.x * 32 + threadIdx.x;
  if (i >= p.n) return;
  const int t = p.tree0 + i, T = p.T, n_ops = T - 2;
  int2* __restrict__ ch = reinterpret_cast<int2*>(p.children) + (size_t)t * n_ops;
  double2* __restrict__ bl = reinterpret_cast<double2*>(p.blen) + (size_t)t * n_ops;
  const char* text = p.text + __ldg(p.offs + i);
  Reader r(text);

  int depth = 0, emitted = 0, cur_n = 0;
  int reg_id = 0, root_child = 0, top0 = 0, top1 = 0;
  double reg_len = 0.0, root_blen = 0.0, topl0 = 0.0, topl1 = 0.0;
  bool ok = true, done = false;
  // a finished child (tip or inner node) joins the node that is open at `depth`
  auto attach = [&](int id, double len) {
    if (depth == 1) {
      if (cur_n == 0) { root_child = id; root_blen = len; }
      else if (cur_n == 1) { top0 = id; topl0 = len; }
      else if (cur_n == 2) { top1 = id; topl1 = len; }
      else ok = false;
    } else {
      if (cur_n == 0) {
        reinterpret_cast<int*>(ch + (n_ops - depth))[0] = id;
        reinterpret_cast<double*>(bl + (n_ops - depth))[0] = len;
      } else if (cur_n == 1) {
        reg_id = id;
        reg_len = len;
      } else {
        ok = false;
      }
    }
    ++cur_n;
  };

  skip(r);
  if (r.peek() != '(') ok = false;
  while (ok && !done) {
    skip(r);
    const int c = r.peek();
    if (c == '(') {
      r.adv();
      ++depth;
      if (emitted + depth > n_ops) { ok = false; break; }  // more open nodes than operations left
      if (depth >= 2) reinterpret_cast<int*>(ch + (n_ops - depth))[1] = cur_n;  // restored at its ')'
      cur_n = 0;
    } else if (c == ',') {
      r.adv();
    } else if (c == ')') {
      r.adv();
      if (depth == 0) { ok = false; break; }
      skip(r);
      while (!is_delim(r.peek())) r.adv();  // inner label, ignored
      double len;
      if (!branch_length(r, &len)) { ok = false; break; }
      if (depth == 1) {
        // the trifurcation (a trailing ":len" on it is accepted and ignored)
        skip(r);
        if (r.peek() != ';' || cur_n != 3 || emitted != n_ops - 1) { ok = false; break; }
        ++emitted;
        done = true;
      } else {
        if (cur_n != 2 || emitted >= n_ops - 1) { ok = false; break; }
        const int2 s0 = ch[n_ops - depth];
        const double l0 = reinterpret_cast<const double*>(bl + (n_ops - depth))[0];
        ch[emitted] = make_int2(s0.x, reg_id);
        bl[emitted] = make_double2(l0, reg_len);
        --depth;
        cur_n = s0.y;
        attach(T + emitted, len);
        ++emitted;
      }
    } else if (c == 0 || c == ';' || c == ':') {
      ok = false;
    } else {
      const uint32_t start = r.pos;
      uint32_t h = 2166136261u;
      uint64_t first8 = 0;
      int len = 0;
      for (int cc = c; !is_delim(cc); cc = r.peek()) {
        h = (h ^ (uint32_t)cc) * 16777619u;
        if (len < 8) first8 |= (uint64_t)cc << (8 * len);
        ++len;
        r.adv();
      }
      if (depth == 0) { ok = false; break; }
      const int row = find_label(p.labels, text, start, len, h, first8);
      double blen_v;
      if (row < 0 || !branch_length(r, &blen_v)) { ok = false; break; }
      attach(row, blen_v);
    }
  }
  if (ok) {
    ch[n_ops - 1] = make_int2(top0, top1);
    bl[n_ops - 1] = make_double2(topl0, topl1);
    p.root_child[t] = root_child;
    p.root_blen[t] = root_blen;
    p.on_pat[t] = (root_child == 0 || top0 == 0 || top1 == 0) ? 1 : 0;
    p.status[t] = 0;
  } else {
    // stand-in: a caterpillar over tips 1..T-1 with the naive tip on the root edge (valid for k_plan);
    // the host parser decides what the tree really is and its result replaces this one
    for (int op = 0; op < n_ops; ++op) {
      ch[op] = op == 0 ? make_int2(1, 2) : make_int2(T + op - 1, op + 2);
      bl[op] = make_double2(kEpsLen, kEpsLen);
    }
    p.root_child[t] = 0;
    p.root_blen[t] = kEpsLen;
    p.on_pat[t] = 1;
    p.status[t] = 1;
  }
}

__global__ void __launch_bounds__(32) k_split(const uint8_t* __restrict__ on_pat, int t0, int n, int32_t* __restrict__ list,
                                              int32_t* __restrict__ n_pat_out) {
  const int lane = threadIdx.x;
  int w = 0;
  for (int pass = 1; pass >= 0; --pass) {
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      const bool mine = i < n && (on_pat[t0 + i] != 0) == (pass == 1);
      const unsigned m = __ballot_sync(0xffffffffu, mine);
      if (mine) list[t0 + w + __popc(m & ((1u << lane) - 1u))] = t0 + i;
      w += __popc(m);
    }
    if (pass == 1 && lane == 0) *n_pat_out = w;
  }
}

}  // namespace

void launch_newick(const NewickDev& p, cudaStream_t s) {
  if (p.n <= 0) return;
  k_newick<<<(p.n + 31) / 32, 32, 0, s>>>(p);
}

void launch_split(const uint8_t* on_pat, int t0, int n, int32_t* list, int32_t* n_pat_out, cudaStream_t s) {
  k_split<<<1, 32, 0, s>>>(on_pat, t0, n, list, n_pat_out);
}

}  // namespace lhb
