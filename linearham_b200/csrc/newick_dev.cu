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

#include <algorithm>
#include <cstdint>
#include <mutex>

#include "decimal.hpp"
#include "kernels.hpp"

namespace lhb {
namespace {

constexpr double kEpsLen = 1e-6;  // src/utils.hpp:20

struct Reader {
  const char* s;
  uint64_t cur;
  uint32_t pos;
  // str is 8-byte aligned; reading starts at character p0
  __device__ __forceinline__ explicit Reader(const char* str, uint32_t p0 = 0) : s(str), pos(p0) {
    cur = __ldg(reinterpret_cast<const unsigned long long*>(s + (p0 & ~7u)));
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
  if (p.only_deferred && p.status[t] != 2) return;   // second pass behind k_newick_cta: the trees it found too deep
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

// ------------------------------------------------------------------ k_newick_cta ---
// The same parse, one CTA per tree, for the latency of a call: a 1001-tip string is 73 KB and a single thread
// needs tens of milliseconds for it (a dependent chain over the characters), which is half a C4 step in front of
// the first pruning kernel. Here the string is cut into one slice per thread and lexed in parallel:
//   S0  brackets: a character is inside a comment iff the nearest bracket to its left is '[' - each thread finds
//       the last bracket of its slice, a warp ballot / per-warp summary gives every slice its state at its start;
//   S1  every thread lexes its slice without knowing what came before it: the state (inside a word?, last token)
//       is fixed from the first structural character on, so a slice is summarised by what it holds before that
//       character (a word run that may continue the previous slice, at most one new word) and by its counts
//       after it; thread 0 folds the summaries front to back (one short loop) into the exact state - last token,
//       depth, ')' count, node count - at the start of every slice;
//   S3  every thread lexes its slice again, now exactly: checks the token grammar and records one item per node
//       (tip: where its label starts; ')': where it ends, which operation it is) with its depth, in text order;
//   S5  one thread per node: label -> xMSA row (hash table), branch length -> double (Eisel-Lemire);
//   S6  children to parents, in parallel (link_items below): a tip or a finished node is filed under its level, a ')'
//       takes the two filed under its level as its children - the stack of the sequential parser, indexed by the
//       depth S3 knows, and resolved with stable ranks per level instead of a walk.
// The grammar accepted here is the strict one - '(' item ',' item [',' item] ')' [label] [':' number], ';' at depth
// 0, nothing after it; every other spelling the lenient host parser takes (missing commas, text after ';', ...)
// is left to it, like everything else this kernel does not decide.
enum : int { K_NONE = 0, K_OPEN, K_CLOSE, K_COMMA, K_COLON, K_SEMI, K_TIP, K_NUM, K_INNER, K_BAD };

// The kind of a word follows from the token before it, and which tokens may come before a structural token is a
// 5 x 10 table; both are looked up for every character, as shifts of 64-bit constants (written as chained
// comparisons they were 15 % of the kernel's instructions).
//   word after '(' or ','  -> tip label;  after ':' -> number;  after ')' -> inner label;  anything else -> K_BAD
__device__ __forceinline__ int word_type(int last) {
  constexpr unsigned long long tbl = (unsigned long long)K_BAD * 0x1111111111ull                       // ten entries of four bits
      ^ ((unsigned long long)(K_BAD ^ K_TIP) << (4 * K_OPEN)) ^ ((unsigned long long)(K_BAD ^ K_TIP) << (4 * K_COMMA))
      ^ ((unsigned long long)(K_BAD ^ K_NUM) << (4 * K_COLON)) ^ ((unsigned long long)(K_BAD ^ K_INNER) << (4 * K_CLOSE));
  return (int)((tbl >> (4 * last)) & 15ull);
}
//   '(' after nothing, '(' or ',';  ')' and ',' after a node (tip, number, ')', inner label);  ':' after a tip, an
//   inner label or ')';  ';' after ')', an inner label or a number;  nothing after ';'
__device__ __forceinline__ bool may_follow(int prev, int cur) {   // cur: K_OPEN..K_SEMI
  constexpr unsigned long long after_node = (1u << K_TIP) | (1u << K_NUM) | (1u << K_CLOSE) | (1u << K_INNER);
  constexpr unsigned long long tbl =
      ((unsigned long long)((1u << K_NONE) | (1u << K_OPEN) | (1u << K_COMMA)) << (10 * (K_OPEN - 1))) |
      (after_node << (10 * (K_CLOSE - 1))) | (after_node << (10 * (K_COMMA - 1))) |
      ((unsigned long long)((1u << K_TIP) | (1u << K_INNER) | (1u << K_CLOSE)) << (10 * (K_COLON - 1))) |
      ((unsigned long long)((1u << K_CLOSE) | (1u << K_INNER) | (1u << K_NUM)) << (10 * (K_SEMI - 1)));
  return (tbl >> (10 * (cur - 1) + prev)) & 1ull;
}
__device__ __forceinline__ int structural(int c) {
  return c == '(' ? K_OPEN : c == ')' ? K_CLOSE : c == ',' ? K_COMMA : c == ':' ? K_COLON : c == ';' ? K_SEMI : K_NONE;
}
__device__ __forceinline__ bool is_blank(int c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

struct SliceSummary {        // what a slice says without knowing what precedes it (S1)
  uint8_t lead_word;         // starts (outside a comment) with a word character: may continue the previous slice's word
  uint8_t early_runs;        // word runs that start before its first structural token (0, 1, 2 = more)
  uint8_t first_tok;         // its first structural token, K_NONE if it has none
  uint8_t last_tok;          // its last token if that is fixed inside the slice, else K_NONE (= the early word, or nothing)
  uint8_t in_word_out;       // its last character is a word character outside a comment
  uint8_t nonempty;
  uint16_t n_open, n_close, n_tips;  // after (and including) the first structural token
};
struct SliceStart {          // the exact lexer state at the start of a slice (thread 0's fold)
  uint8_t in_word, last_tok;
  uint16_t depth;
  int32_t n_close, n_items;
};

constexpr int kCtaLevels = 256;   // depth the per-level tables are sized for; deeper trees go to k_newick
// Per-tree item arrays (global scratch, L2-resident: written in S3/S5, read in S5/S6 by the same CTA):
//   len [N] double (before S5: the text position in the low word) | id [N] int32 (xMSA row of a tip /
//   operation index of a ')') | lvl [N] uint16 (depth of the parent's child list; bit 15: the node is a ')')
//   rnk [N] uint16 (S6: the node's rank among the nodes filed under its level) | tab [T-2] int32 (S6: the ')' items
//   grouped by level, in text order)
__host__ __device__ inline size_t cta_items_stride(int n_tips) {
  return ((size_t)(2 * n_tips - 2) * 16 + (size_t)(n_tips - 2) * 4 + 15) / 16 * 16;
}
struct LevelTables {         // S6, per level: ')' of that level, where its group starts in tab, running ranks
  int n_close[kCtaLevels + 1], start[kCtaLevels + 1], run_filed[kCtaLevels + 1], run_close[kCtaLevels + 1];
};
struct CtaLayout {           // dynamic shared memory of k_newick_cta
  size_t lev, sum, start, total;
  __host__ __device__ explicit CtaLayout(int nt) {
    size_t at = 0;
    auto take = [&](size_t bytes) { const size_t o = at; at += (bytes + 15) / 16 * 16; return o; };
    lev = take(sizeof(LevelTables));
    sum = take((size_t)nt * sizeof(SliceSummary));
    start = take((size_t)nt * sizeof(SliceStart));
    total = at;
  }
};

// ---- S6: children to parents, in parallel -------------------------------------------------------------------------
// A tip is filed under the level of the list it stands in, a ')' under the level below the list it closes, and a ')'
// takes the two nodes filed under its own level since the previous ')' of that level: per level the text reads
// (F F C)*. So the k-th ')' of level l (in text order) owns the nodes of rank 2k and 2k+1 among those filed under l.
// The ranks are stable counting-sort ranks: one warp walks the items in text order, 32 at a time, and ranks a window
// with __match_any_sync plus a running count per level; the ')' items are scattered into `tab`, grouped by level.
// Then every node finds its parent on its own - rank >> 1 names the ')' - and checks that it stands between that ')'
// and the one before it; with "filed = 2 x closes" per level that is exactly the (F F C)* pattern. Level 1 is the
// root's list: three nodes (root_child, then the two kept in the last operation) and one ')' that must be the last
// item. NT threads (a CTA, or one warp) call this together; returns false (to every thread) if the tree is not of
// that shape. The serial walk this replaces took ~0.8 ms per 1001-tip tree (2000 dependent steps over L2-resident
// items): the latency of a small chunk's parse, which is what stands in front of its pruning kernel.
template <int NT>
__device__ __forceinline__ void group_sync() {
  if (NT == 32) __syncwarp();
  else __syncthreads();
}
template <int NT>
__device__ bool link_items(const NewickDev& p, int t, int T, int n_items, const double* LEN, const int32_t* ID,
                           const uint16_t* LVL, uint16_t* RNK, int32_t* TAB, LevelTables& lev, int* s_bad, int* s_on_pat,
                           int2* __restrict__ ch, double2* __restrict__ bl) {
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, n_ops = T - 2;
  for (int l = tid; l <= kCtaLevels; l += NT) lev.n_close[l] = lev.run_filed[l] = lev.run_close[l] = 0;
  if (tid == 0) *s_bad = *s_on_pat = 0;
  group_sync<NT>();
  for (int it = tid; it < n_items; it += NT) {
    const unsigned lv = __ldcg(LVL + it);
    if (lv & 0x8000u) atomicAdd(&lev.n_close[lv & 0x7fffu], 1);
  }
  group_sync<NT>();
  if (tid < 32) {   // exclusive scan over the levels
    int carry = 0;
    for (int base = 0; base <= kCtaLevels; base += 32) {
      const int l = base + lane, v = l <= kCtaLevels ? lev.n_close[l] : 0;
      int inc = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(FULL, inc, d);
        if (lane >= d) inc += up;
      }
      if (l <= kCtaLevels) lev.start[l] = carry + inc - v;
      carry += __shfl_sync(FULL, inc, 31);
    }
    __syncwarp();
    // stable ranks, in text order
    const unsigned lt = (1u << lane) - 1u;
    for (int base = 0; base < n_items; base += 32) {
      const int it = base + lane;
      const bool valid = it < n_items;
      const unsigned lv = valid ? __ldcg(LVL + it) : 0u;
      const bool close = (lv & 0x8000u) != 0;
      const int lvl = (int)(lv & 0x7fffu);
      const int filed = valid ? (close ? lvl - 1 : lvl) : 0x10000 + lane;     // (an idle lane groups with nobody)
      const int closed = (valid && close) ? lvl : 0x10000 + lane;
      const unsigned mf = __match_any_sync(FULL, filed), mc = __match_any_sync(FULL, closed);
      int rf = 0, rc = 0;
      if (valid) rf = lev.run_filed[filed] + __popc(mf & lt);
      if (valid && close) rc = lev.run_close[lvl] + __popc(mc & lt);
      __syncwarp();
      if (valid && lane == 31 - __clz(mf)) lev.run_filed[filed] += __popc(mf);
      if (valid && close && lane == 31 - __clz(mc)) lev.run_close[lvl] += __popc(mc);
      __syncwarp();
      if (valid) RNK[it] = (uint16_t)rf;
      if (valid && close) TAB[lev.start[lvl] + rc] = it;
    }
  }
  group_sync<NT>();
  bool bad = false;
  for (int l = tid; l <= kCtaLevels; l += NT) {
    if (l >= 2 && lev.run_filed[l] != 2 * lev.n_close[l]) bad = true;
    if (l == 1 && (lev.run_filed[1] != 3 || lev.n_close[1] != 1)) bad = true;
  }
  for (int it = tid; it < n_items; it += NT) {
    const unsigned lv = __ldcg(LVL + it);
    const bool close = (lv & 0x8000u) != 0;
    const int lvl = (int)(lv & 0x7fffu), filed = close ? lvl - 1 : lvl, r = __ldcg(RNK + it);
    const int own = __ldcg(ID + it), id = close ? T + own : own;
    const double len = __ldcg(LEN + it);
    if (filed == 0) {   // the ')' of the root's list: the last item, the last operation
      if (it != n_items - 1 || own != n_ops - 1) bad = true;
      continue;
    }
    if (filed == 1) {
      if (r == 0) {
        p.root_child[t] = id;
        p.root_blen[t] = len;
      } else if (r == 1) {
        ch[n_ops - 1].x = id;
        bl[n_ops - 1].x = len;
      } else if (r == 2) {
        ch[n_ops - 1].y = id;
        bl[n_ops - 1].y = len;
      } else {
        bad = true;
      }
      if (id == 0) *s_on_pat = 1;   // the naive tip under the root: the tree runs on site patterns
      continue;
    }
    const int k = r >> 1;
    if (k >= lev.n_close[filed]) { bad = true; continue; }
    const int32_t* grp = TAB + lev.start[filed];
    const int parent = __ldcg(grp + k);
    if (it > parent || (k > 0 && it < __ldcg(grp + k - 1))) { bad = true; continue; }
    const int op = __ldcg(ID + parent);
    if (r & 1) {
      ch[op].y = id;
      bl[op].y = len;
    } else {
      ch[op].x = id;
      bl[op].x = len;
    }
  }
  if (bad) *s_bad = 1;
  group_sync<NT>();
  return *s_bad == 0;
}

// (at most 40 registers - 22 bytes of spills, 14.3 ms instead of 14.8 per 16 384 C4 trees. 128 x 40 = 5 K registers
// is also what one k_prune CTA of 32 x 164 leaves free when it steps aside - option "parse_reserve" - so that the
// parser of the next chunk can run beside the pruning of this one. Measured (r4i, C4): that buys nothing, 160 k
// trees/s with 0 CTAs reserved, 157 k with 1, 150 k with 2 - the GPU is busy either way and the parser's SM time is
// paid beside k_prune or between its launches - so the option stays 0.)
template <int NT>
__global__ void __launch_bounds__(NT, 12) k_newick_cta(NewickDev p) {
  extern __shared__ __align__(16) unsigned char nw_smem[];
  // s_flag[j] != 0: stage j (0 fold, 1 S3, 2 S5, 3 S6) leaves the tree to the host. One flag per stage: a stage reads
  // only flags written before the barrier in front of it, so every thread takes the same branch around the barriers.
  __shared__ int s_flag[4], s_deep, s_warp_has[NT / 32], s_warp_open[NT / 32];
  const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int t = p.tree0 + i, T = p.T, n_ops = T - 2;
  const CtaLayout lay(NT);
  const int n_items = 2 * T - 2;   // T tips + T-2 ')'
  unsigned char* items = p.items + (size_t)i * cta_items_stride(T);
  double* LEN = reinterpret_cast<double*>(items);
  int32_t* ID = reinterpret_cast<int32_t*>(items + (size_t)n_items * 8);
  uint16_t* LVL = reinterpret_cast<uint16_t*>(items + (size_t)n_items * 12);
  uint16_t* RNK = reinterpret_cast<uint16_t*>(items + (size_t)n_items * 14);
  int32_t* TAB = reinterpret_cast<int32_t*>(items + (size_t)n_items * 16);
  LevelTables& lev = *reinterpret_cast<LevelTables*>(nw_smem + lay.lev);
  SliceSummary* SUM = reinterpret_cast<SliceSummary*>(nw_smem + lay.sum);
  SliceStart* START = reinterpret_cast<SliceStart*>(nw_smem + lay.start);
  int2* __restrict__ ch = reinterpret_cast<int2*>(p.children) + (size_t)t * n_ops;
  double2* __restrict__ bl = reinterpret_cast<double2*>(p.blen) + (size_t)t * n_ops;
  const char* text = p.text + __ldg(p.offs + i);
  const uint32_t L = __ldg(p.lens + i);
  const uint32_t slice = ((L + NT - 1) / NT + 7u) & ~7u;          // a multiple of 8: slices start on the reader's words
  const uint32_t a = min(L, (uint32_t)tid * slice), b = min(L, a + slice);
  // character classes, one shared-memory look-up per character instead of a dozen compares: 0 word character,
  // K_OPEN..K_SEMI the structural ones, 6 '[', 7 ']' (a word character unless it closes a comment), 8 blank / NUL
  __shared__ uint8_t s_cls[256];
  for (int c = tid; c < 256; c += NT)
    s_cls[c] = (uint8_t)(c == '[' ? 6 : c == ']' ? 7 : (c == 0 || is_blank(c)) ? 8 : structural(c));
  if (tid < 4) s_flag[tid] = 0;
  if (tid == 4) s_deep = 0;

  // ---- S0: is the start of my slice inside a comment? ----
  int my_last = 0;  // 0: no bracket in the slice, 1: its last bracket is '[', 2: ']'
  for (uint32_t q = a; q < b; q += 8) {
    const unsigned long long v = __ldg(reinterpret_cast<const unsigned long long*>(text + q));
    const int n = (int)min(8u, b - q);
    for (int j = 0; j < n; ++j) {
      const int c = (int)((v >> (8 * j)) & 0xff);
      if (c == '[') my_last = 1;
      else if (c == ']') my_last = 2;
    }
  }
  const unsigned has = __ballot_sync(0xffffffffu, my_last != 0), open = __ballot_sync(0xffffffffu, my_last == 1);
  if (lane == 0) {
    s_warp_has[wid] = has != 0;
    s_warp_open[wid] = has ? (int)((open >> (31 - __clz(has))) & 1u) : 0;
  }
  __syncthreads();
  bool in_comment = false;
  {
    const unsigned before = has & ((1u << lane) - 1u);
    if (before) {
      in_comment = (open >> (31 - __clz(before))) & 1u;
    } else {
      for (int w = wid - 1; w >= 0; --w)
        if (s_warp_has[w]) {
          in_comment = s_warp_open[w] != 0;
          break;
        }
    }
  }

  // ---- S1: what my slice holds, not knowing what precedes it ----
  {
    SliceSummary sm{};
    sm.nonempty = b > a;
    bool com = in_comment, in_word = false, synced = false;
    int last = K_NONE;  // the last token since the first structural one
    Reader r(text, a);
    for (uint32_t q = a; q < b; ++q, r.adv()) {
      const int cls = s_cls[r.peek()];
      if (com) {
        com = cls != 7;
        continue;
      }
      if (cls >= 6 && cls != 7) {  // '[' or a blank: either ends a word
        com = cls == 6;
        in_word = false;
        continue;
      }
      const int st = cls == 7 ? K_NONE : cls;
      if (st != K_NONE) {
        if (!synced) { synced = true; sm.first_tok = (uint8_t)st; }
        sm.n_open += st == K_OPEN;
        sm.n_close += st == K_CLOSE;
        last = st;
        in_word = false;
        continue;
      }
      if (!in_word) {  // a word starts here
        in_word = true;
        if (!synced) {
          if (q == a) sm.lead_word = 1;
          sm.early_runs = sm.early_runs < 2 ? sm.early_runs + 1 : 2;
        } else {
          const int wt = word_type(last);
          sm.n_tips += wt == K_TIP;
          last = wt;
        }
      }
    }
    sm.last_tok = (uint8_t)(synced ? last : K_NONE);
    sm.in_word_out = in_word && !com;
    SUM[tid] = sm;
  }
  __syncthreads();

  // ---- thread 0 folds the summaries into the exact state at the start of every slice ----
  if (tid == 0) {
    int in_word = 0, last = K_NONE, depth = 0, n_close = 0, n_items = 0;
    for (int k = 0; k < NT; ++k) {
      SliceStart st;
      st.in_word = (uint8_t)in_word;
      st.last_tok = (uint8_t)last;
      st.depth = depth;
      st.n_close = n_close;
      st.n_items = n_items;
      START[k] = st;
      const SliceSummary sm = SUM[k];
      if (!sm.nonempty) continue;
      const int runs = (int)sm.early_runs - ((sm.lead_word && in_word) ? 1 : 0);
      if (runs >= 1) {  // a new word before the slice's first structural token: its kind follows from what came before
        const int wt = word_type(last);
        n_items += wt == K_TIP;
        last = wt;
      }
      if (sm.first_tok != K_NONE) {
        depth += (int)sm.n_open - (int)sm.n_close;
        n_items += (int)sm.n_tips + (int)sm.n_close;
        n_close += sm.n_close;
        last = sm.last_tok;
      }
      in_word = sm.in_word_out;
    }
    // a tree of this family ends with ';' at depth 0 and has T tips and T-2 ')'
    if (last != K_SEMI || depth != 0 || n_close != n_ops || n_items != 2 * T - 2) s_flag[0] = 1;
  }
  __syncthreads();

  if (!s_flag[0]) {
    // ---- S3: the exact pass: token grammar, one item per node ----
    {
      const SliceStart st = START[tid];
      bool com = in_comment, in_word = st.in_word != 0, bad = false;
      int last = st.last_tok, depth = st.depth, nc = st.n_close, ni = st.n_items;
      Reader r(text, a);
      for (uint32_t q = a; q < b; ++q, r.adv()) {
        const int cls = s_cls[r.peek()];
        if (com) {
          com = cls != 7;
          continue;
        }
        if (cls >= 6 && cls != 7) {
          com = cls == 6;
          in_word = false;
          continue;
        }
        const int stc = cls == 7 ? K_NONE : cls;
        if (stc != K_NONE) {
          in_word = false;
          if (!may_follow(last, stc)) bad = true;   // (nothing may follow ';')
          if (stc == K_OPEN) {
            if (++depth > kCtaLevels) { bad = true; s_deep = 1; }
          } else if (stc == K_CLOSE) {
            if (depth <= 0 || ni >= n_items) {
              bad = true;
            } else {
              LVL[ni] = (uint16_t)(depth | 0x8000);
              ID[ni] = nc;
              reinterpret_cast<uint32_t*>(LEN + ni)[0] = q + 1;
            }
            ++ni;
            ++nc;
            --depth;
          } else if (stc == K_COMMA) {
            if (depth <= 0) bad = true;
          } else if (stc == K_SEMI) {
            if (depth != 0) bad = true;
          }
          last = stc;
          continue;
        }
        if (!in_word) {
          in_word = true;
          const int wt = word_type(last);
          if (wt == K_BAD) bad = true;
          if (wt == K_TIP) {
            if (depth <= 0 || ni >= n_items) {
              bad = true;
            } else {
              LVL[ni] = (uint16_t)depth;
              reinterpret_cast<uint32_t*>(LEN + ni)[0] = q;
            }
            ++ni;
          }
          last = wt;
        }
      }
      if (bad) s_flag[1] = 1;
    }
    __syncthreads();
  }

  if (!s_flag[0] && !s_flag[1]) {
    // ---- S5: one thread per node: label -> row, branch length -> double ----
    bool bad = false;
    for (int it = tid; it < n_items; it += NT) {
      const uint32_t pos = __ldcg(reinterpret_cast<const uint32_t*>(LEN + it));
      Reader r(text, pos);
      if ((__ldcg(LVL + it) & 0x8000u) == 0) {
        uint32_t h = 2166136261u;
        uint64_t first8 = 0;
        int len = 0;
        for (int cc = r.peek(); !is_delim(cc); cc = r.peek()) {
          h = (h ^ (uint32_t)cc) * 16777619u;
          if (len < 8) first8 |= (uint64_t)cc << (8 * len);
          ++len;
          r.adv();
        }
        const int row = find_label(p.labels, text, pos, len, h, first8);
        if (row < 0) bad = true;
        ID[it] = row;
      } else {
        skip(r);
        while (!is_delim(r.peek())) r.adv();  // inner label, ignored
      }
      double v = 0.0;
      if (!branch_length(r, &v)) bad = true;
      LEN[it] = v;
    }
    if (bad) s_flag[2] = 1;
    __syncthreads();
  }

  if (!s_flag[0] && !s_flag[1] && !s_flag[2]) {
    // ---- S6: children to parents ----
    __shared__ int s_bad, s_on_pat;
    const bool ok = link_items<NT>(p, t, T, n_items, LEN, ID, LVL, RNK, TAB, lev, &s_bad, &s_on_pat, ch, bl);
    if (tid == 0) {
      if (ok) {
        p.on_pat[t] = s_on_pat ? 1 : 0;
        p.status[t] = 0;
      } else {
        s_flag[3] = 1;
      }
    }
  }
  __syncthreads();
  if (s_flag[0] | s_flag[1] | s_flag[2] | s_flag[3]) {
    // stand-in (see k_newick): the host parser decides what the tree really is
    for (int op = tid; op < n_ops; op += NT) {
      ch[op] = op == 0 ? make_int2(1, 2) : make_int2(T + op - 1, op + 2);
      bl[op] = make_double2(kEpsLen, kEpsLen);
    }
    if (tid == 0) {
      p.root_child[t] = 0;
      p.root_blen[t] = kEpsLen;
      p.on_pat[t] = 1;
      p.status[t] = s_deep ? 2 : 1;   // 2: deeper than this kernel's tables - k_newick takes the tree
    }
  }
}
// ------------------------------------------------------------------ k_newick_warp ---
// One warp per tree, the lexing data-parallel over the characters instead of over slices: the warp reads 32
// consecutive characters per step (one 32-byte sector, coalesced) and every lane classifies its own; what a
// sequential lexer carries from character to character becomes warp-wide bit arithmetic on ballot masks -
//   comments   the bracket masks are walked once per window (a loop over their set bits, the same in every lane):
//              '[' outside a comment opens one, the next ']' closes it, ']' outside one is a word character;
//   words      W = word characters outside comments; a word starts where W has a bit and its left neighbour has not;
//   tokens     T = structural characters | word starts; the kind of a word follows from the token before it (its
//              lane is the highest bit of T below mine, read with one shuffle), the grammar check from that token's
//              kind (a second shuffle);
//   counters   depth, node index and ')' index of a lane = the carried value + popcounts of the masks below it.
// Every lane is busy on every step and no lane diverges, so a 73 KB string costs ~2300 steps of ~70 instructions
// instead of two divergent byte loops per thread, and there is nothing to summarise and fold: the state at the end
// of a window (in a comment?, in a word?, last token, depth, counts) is carried in registers to the next. S5 (label
// -> row, branch length -> double, one lane per node) and S6 (children to parents) are k_newick_cta's.
__global__ void __launch_bounds__(32, 24) k_newick_warp(NewickDev p) {
  __shared__ LevelTables lev;
  __shared__ int s_bad, s_on_pat;
  __shared__ uint8_t s_cls[256];
  constexpr unsigned FULL = 0xffffffffu;
  const int i = blockIdx.x, lane = threadIdx.x;
  const int t = p.tree0 + i, T = p.T, n_ops = T - 2;
  const int n_items = 2 * T - 2;   // T tips + T-2 ')'
  unsigned char* items = p.items + (size_t)i * cta_items_stride(T);
  double* LEN = reinterpret_cast<double*>(items);
  int32_t* ID = reinterpret_cast<int32_t*>(items + (size_t)n_items * 8);
  uint16_t* LVL = reinterpret_cast<uint16_t*>(items + (size_t)n_items * 12);
  uint16_t* RNK = reinterpret_cast<uint16_t*>(items + (size_t)n_items * 14);
  int32_t* TAB = reinterpret_cast<int32_t*>(items + (size_t)n_items * 16);
  int2* __restrict__ ch = reinterpret_cast<int2*>(p.children) + (size_t)t * n_ops;
  double2* __restrict__ bl = reinterpret_cast<double2*>(p.blen) + (size_t)t * n_ops;
  const char* text = p.text + __ldg(p.offs + i);
  const uint32_t L = __ldg(p.lens + i);
  for (int c = lane; c < 256; c += 32)
    s_cls[c] = (uint8_t)(c == '[' ? 6 : c == ']' ? 7 : (c == 0 || is_blank(c)) ? 8 : structural(c));
  __syncwarp();

  // ---- lexing: token grammar, one item per node (what S0-S3 of k_newick_cta do in four passes) ----
  bool bad = false, deep = false;
  {
    const unsigned lt = (1u << lane) - 1u;
    unsigned com = 0;              // the window starts inside a comment
    unsigned in_word = 0;          // the character before the window is a word character
    int last = K_NONE, depth = 0, ni = 0, nc = 0;
    const unsigned char* tx = reinterpret_cast<const unsigned char*>(text);
    int c_next = lane < L ? (int)__ldg(tx + lane) : ' ';
    for (uint32_t base = 0; base < L; base += 32) {
      const uint32_t q = base + lane;
      const int cls = s_cls[c_next];
      c_next = q + 32 < L ? (int)__ldg(tx + q + 32) : ' ';   // (the next window's load is in flight during this one)
      // the four bits of every lane's class: all class masks follow from them, the same in every lane
      const unsigned b0 = __ballot_sync(FULL, cls & 1), b1 = __ballot_sync(FULL, cls & 2);
      const unsigned b2 = __ballot_sync(FULL, cls & 4), b3 = __ballot_sync(FULL, cls & 8);
      const unsigned m_ob = b1 & b2 & ~b0, m_cb = b0 & b1 & b2;          // '[' (6), ']' (7)
      // Comments. After a '[' the text is inside a comment whatever it was before, after a ']' outside one: a
      // character is inside iff the nearest bracket to its left is '['. Adding a one above every '[' to the mask of
      // non-brackets lets the carry run up to the next bracket; the bits the addition changes are that stretch.
      const unsigned nb = ~(m_ob | m_cb);
      const unsigned com_mask = ((nb + ((m_ob << 1) | com)) ^ nb) | m_ob;
      if (~nb) com = (m_ob >> (31 - __clz(~nb))) & 1u;
      const unsigned U = (b0 | b1 | b2) & ~(b1 & b2) & ~com_mask;         // structural characters (1..5)
      const unsigned W = (~(b0 | b1 | b2 | b3) | m_cb) & ~com_mask;       // word characters (0; ']' outside a comment)
      const unsigned MO = b0 & ~b1 & ~b2 & ~com_mask, MC = b1 & ~b0 & ~b2 & ~com_mask;   // '(' (1), ')' (2)
      const unsigned WS = W & ~((W << 1) | in_word);
      const unsigned TK = U | WS;
      const int st = ((U >> lane) & 1u) ? cls : K_NONE;
      const unsigned below = TK & lt;
      const int pidx = below ? 31 - __clz(below) : 0;
      const int st_prev = __shfl_sync(FULL, st, pidx);
      const bool word_start = (WS >> lane) & 1u;
      int tok = st;
      if (word_start) tok = below ? word_type(st_prev) : word_type(last);   // (after a word: K_BAD)
      const int tok_prev = __shfl_sync(FULL, tok, pidx);
      const int prev = below ? tok_prev : last;
      const bool is_tip = word_start && tok == K_TIP;
      const unsigned IT = __ballot_sync(FULL, is_tip) | MC;
      const int d = depth + __popc(MO & lt) - __popc(MC & lt);   // the depth in front of my character
      const int my = ni + __popc(IT & lt);
      if (st != K_NONE) {
        if (!may_follow(prev, st)) bad = true;   // (nothing may follow ';')
        if (st == K_OPEN) {
          if (d + 1 > kCtaLevels) { bad = true; deep = true; }
        } else if (st == K_SEMI) {
          if (d != 0) bad = true;
        } else if (st != K_COLON && d <= 0) {           // ')' and ','
          bad = true;
        }
      } else if (word_start && tok == K_BAD) {
        bad = true;
      }
      if ((IT >> lane) & 1u) {
        if (d <= 0 || my >= n_items) {
          bad = true;
        } else {
          LVL[my] = (uint16_t)(is_tip ? d : (d | 0x8000));
          if (!is_tip) ID[my] = nc + __popc(MC & lt);
          reinterpret_cast<uint32_t*>(LEN + my)[0] = is_tip ? q : q + 1;
        }
      }
      depth += __popc(MO) - __popc(MC);
      ni += __popc(IT);
      nc += __popc(MC);
      in_word = W >> 31;
      if (TK) last = __shfl_sync(FULL, tok, 31 - __clz(TK));
    }
    // a tree of this family ends with ';' at depth 0 and has T tips and T-2 ')'
    if (last != K_SEMI || depth != 0 || nc != n_ops || ni != n_items) bad = true;
    bad = __any_sync(FULL, bad);
    deep = __any_sync(FULL, deep);
  }
  __syncwarp();

  if (!bad) {
    // ---- S5: one lane per node: label -> row, branch length -> double ----
    for (int it = lane; it < n_items; it += 32) {
      const uint32_t pos = __ldcg(reinterpret_cast<const uint32_t*>(LEN + it));
      Reader r(text, pos);
      if ((__ldcg(LVL + it) & 0x8000u) == 0) {
        uint32_t h = 2166136261u;
        uint64_t first8 = 0;
        int len = 0;
        for (int cc = r.peek(); !is_delim(cc); cc = r.peek()) {
          h = (h ^ (uint32_t)cc) * 16777619u;
          if (len < 8) first8 |= (uint64_t)cc << (8 * len);
          ++len;
          r.adv();
        }
        const int row = find_label(p.labels, text, pos, len, h, first8);
        if (row < 0) bad = true;
        ID[it] = row;
      } else {
        skip(r);
        while (!is_delim(r.peek())) r.adv();  // inner label, ignored
      }
      double v = 0.0;
      if (!branch_length(r, &v)) bad = true;
      LEN[it] = v;
    }
    bad = __any_sync(FULL, bad);
    __syncwarp();
  }

  if (!bad) {
    // ---- S6: children to parents ----
    if (link_items<32>(p, t, T, n_items, LEN, ID, LVL, RNK, TAB, lev, &s_bad, &s_on_pat, ch, bl)) {
      if (lane == 0) {
        p.on_pat[t] = s_on_pat ? 1 : 0;
        p.status[t] = 0;
      }
    } else {
      bad = true;
    }
  }
  if (bad) {
    // stand-in (see k_newick): the host parser decides what the tree really is
    for (int op = lane; op < n_ops; op += 32) {
      ch[op] = op == 0 ? make_int2(1, 2) : make_int2(T + op - 1, op + 2);
      bl[op] = make_double2(kEpsLen, kEpsLen);
    }
    if (lane == 0) {
      p.root_child[t] = 0;
      p.root_blen[t] = kEpsLen;
      p.on_pat[t] = 1;
      p.status[t] = deep ? 2 : 1;   // 2: deeper than this kernel's tables - k_newick takes the tree
    }
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

constexpr int kCtaThreads = 128;
size_t newick_items_bytes(int n_tips, int n_trees) { return cta_items_stride(n_tips) * (size_t)n_trees; }

void launch_newick(const NewickDev& p, cudaStream_t s) {
  if (p.n <= 0) return;
  const bool cta = p.kernel != 1 && p.T < 32768 && p.items != nullptr;
  if (cta && p.kernel == 3) {
    k_newick_warp<<<p.n, 32, 0, s>>>(p);
  } else if (cta) {
    const size_t smem = CtaLayout(kCtaThreads).total;
    k_newick_cta<kCtaThreads><<<p.n, kCtaThreads, smem, s>>>(p);
  }
  NewickDev q = p;
  q.only_deferred = cta ? 1 : 0;
  k_newick<<<(p.n + 31) / 32, 32, 0, s>>>(q);
}

// ---- the no-allocation variant (lh_newick_parse_noalloc; no counterpart in the reference) -----------------------
// k_newick_cta's per-tree item scratch in a file-scope device array instead of a cudaMalloc'ed buffer: every GPU
// that loads the library holds kNoAllocScratchBytes until exit. Every call shares the array; the kernel writes an
// item before it reads it (S3 before S5 / S6), so nothing in it has to start at zero and nothing is cleared - but
// two calls that overlap on one GPU race on it.
static __device__ __align__(256) unsigned char g_newick_static_scratch[kNoAllocScratchBytes];

int newick_noalloc_max_trees(int n_tips) {
  if (n_tips < 3 || n_tips >= 32768) return 0;
  return (int)std::min<size_t>((size_t)1 << 30, (size_t)kNoAllocScratchBytes / cta_items_stride(n_tips));
}

// `scratch` in host code is a host-side stand-in for the array, not its device address: the address is asked for
// once per device, with that device current. Returns nullptr (and the CUDA error through *err) on failure.
static unsigned char* static_scratch_address(int device, cudaError_t* err) {
  static std::mutex mu;
  static unsigned char* addr[64] = {};
  *err = cudaSuccess;
  if (device < 0 || device >= 64) {
    *err = cudaErrorInvalidDevice;
    return nullptr;
  }
  std::lock_guard<std::mutex> g(mu);
  if (!addr[device]) {
    int cur = -1;
    *err = cudaGetDevice(&cur);
    if (*err != cudaSuccess) return nullptr;
    if (cur != device) {
      *err = cudaErrorInvalidDevice;   // the caller makes the device current first
      return nullptr;
    }
    void* ptr = nullptr;
    *err = cudaGetSymbolAddress(&ptr, g_newick_static_scratch);
    if (*err != cudaSuccess) return nullptr;
    addr[device] = static_cast<unsigned char*>(ptr);
  }
  return addr[device];
}

// Launches k_newick_cta (and k_newick behind it for trees nested deeper than its tables) over p.n trees with the
// static scratch; allocates nothing. `device` must be current. 0, or the CUDA error code.
int launch_newick_noalloc(NewickDev p, int device, cudaStream_t s) {
  cudaError_t e;
  unsigned char* scratch = static_scratch_address(device, &e);
  if (!scratch) return (int)e;
  if (p.n <= 0) return 0;
  p.items = scratch;
  p.only_deferred = 0;
  k_newick_cta<kCtaThreads><<<p.n, kCtaThreads, CtaLayout(kCtaThreads).total, s>>>(p);
  p.only_deferred = 1;
  k_newick<<<(p.n + 31) / 32, 32, 0, s>>>(p);
  return (int)cudaGetLastError();
}

void launch_split(const uint8_t* on_pat, int t0, int n, int32_t* list, int32_t* n_pat_out, cudaStream_t s) {
  k_split<<<1, 32, 0, s>>>(on_pat, t0, n, list, n_pat_out);
}

}  // namespace lhb
