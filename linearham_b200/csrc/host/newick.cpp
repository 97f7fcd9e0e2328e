// Host tree ingest: Newick strings -> the flat operation arrays of lh_tree_batch_t.
//
// Replaces, per tree sample, the reference's
//   std::regex_replace("[&index=N]") + pll_utree_parse_newick_string   src/PhyloHMM.cpp:419-421
//   set_missing_branch_length(tree, EPS)                               pll_util.cpp:46-62
//   pll_utree_traverse(POSTORDER) + pll_utree_create_operations         utree.c:284-329, 379-442
// without building a node graph: one pass over the characters, an explicit stack of open
// parentheses, no allocation per node. Conventions kept from the reference:
//   * grammar (parse_utree.y:179-246): top level "(a,b,c)" trifurcation, every other inner node
//     binary, optional inner labels (ignored), optional ":length" (atof; missing -> 0)
//   * the virtual root is the trifurcation; the likelihood is evaluated on the edge to its FIRST
//     listed subtree (pll_util.cpp:368-371, pll_partition.cpp:309-315) -> root_child / root_blen
//   * operations in post-order: subtrees in listed order, the trifurcation last with the second
//     and third subtree as its children
//   * a zero or missing branch length becomes EPS = 1e-6 (src/utils.hpp:20)
// Tip labels are mapped to xMSA rows through the label list of the family (xmsa_labels_,
// src/PhyloHMM.cpp:48-53; Partition::SetTipStates matches by label, pll_partition.cpp:153-194).
#include <atomic>
#include <charconv>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/linearham_b200_host.h"
#include "host_error.hpp"

namespace lhb {
namespace {

constexpr double kEps = 1e-6;  // src/utils.hpp:20

struct Open {        // one unclosed '('
  int32_t n;         // children seen so far
  int32_t id[3];
  double len[3];
};

// Tip label -> xMSA row: open addressing on the FNV-1a hash the scanner computes while it walks
// the label, so a lookup builds no std::string.
struct LabelMap {
  struct Slot {
    const char* s = nullptr;
    uint32_t len = 0, hash = 0;
    int32_t row = -1;
  };
  std::vector<Slot> slots;
  uint32_t mask = 0;
  static uint32_t hash_of(const char* s, size_t n) {
    uint32_t h = 2166136261u;
    for (size_t i = 0; i < n; ++i) h = (h ^ (unsigned char)s[i]) * 16777619u;
    return h;
  }
  void init(int n) {
    size_t cap = 8;
    while (cap < (size_t)n * 2) cap <<= 1;
    slots.assign(cap, Slot{});
    mask = (uint32_t)cap - 1;
  }
  bool insert(const char* s, int32_t row) {  // false: duplicate
    const uint32_t len = (uint32_t)std::strlen(s), h = hash_of(s, len);
    for (uint32_t i = h & mask;; i = (i + 1) & mask) {
      Slot& e = slots[i];
      if (e.row < 0) {
        e = Slot{s, len, h, row};
        return true;
      }
      if (e.hash == h && e.len == len && std::memcmp(e.s, s, len) == 0) return false;
    }
  }
  int32_t find(const char* s, uint32_t len, uint32_t h) const {
    for (uint32_t i = h & mask;; i = (i + 1) & mask) {
      const Slot& e = slots[i];
      if (e.row < 0) return -1;
      if (e.hash == h && e.len == len && std::memcmp(e.s, s, len) == 0) return e.row;
    }
  }
};

// delimiter table: ( ) , : ; [ blank tab newline cr nul
struct DelimTable {
  bool t[256] = {};
  constexpr DelimTable() {
    for (unsigned char c : {'(', ')', ',', ':', ';', '[', ' ', '\t', '\n', '\r', '\0'}) t[c] = true;
  }
};
constexpr DelimTable kDelim{};
inline bool is_delim(char c) { return kDelim.t[(unsigned char)c]; }

// skips blanks and [...] comments (the reference strips "[&index=N]" with a regex)
inline const char* skip(const char* p) {
  for (;;) {
    while (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r') ++p;
    if (*p != '[') return p;
    while (*p != ']' && *p) ++p;  // comments are a dozen characters: a call to strchr costs more
    if (*p) ++p;
  }
}

// Decimal -> double. std::from_chars (libstdc++: Eisel-Lemire with an exact fallback) returns the
// correctly rounded double, as strtod / atof (the reference's lexer path, parse_utree.y) does, at a
// fraction of the cost; forms it does not take (leading '+', hex, locale) go to strtod.
inline const char* parse_double(const char* p, const char* end, double* out) {
  if ((*p >= '0' && *p <= '9') || *p == '-' || *p == '.') {
    const auto r = std::from_chars(p, end, *out);  // stops at the first character that is not part of the number
    if (r.ec == std::errc() && is_delim(*r.ptr)) return r.ptr;
  }
  char* stop = nullptr;
  *out = std::strtod(p, &stop);
  return stop == p ? nullptr : stop;
}

inline const char* branch_length(const char* p, const char* end, double* len) {
  p = skip(p);
  double v = 0.0;
  if (*p == ':') {
    p = parse_double(skip(p + 1), end, &v);
    if (!p) return nullptr;
  }
  *len = v ? v : kEps;
  return p;
}

// Returns nullptr on success, else a static message.
const char* flatten_one(const char* s, const LabelMap& labels, int T, std::vector<Open>& stack,
                        std::vector<uint8_t>& seen, int32_t* children, double* blen, int32_t* root_child,
                        double* root_blen, std::string& detail) {
  const int n_ops = T - 2;
  stack.clear();
  seen.assign(T, 0);
  int emitted = 0;
  bool done = false;
  const char* const end = s + std::strlen(s);
  const char* p = skip(s);
  if (*p != '(') return "tree does not start with '('";
  while (!done) {
    p = skip(p);
    const char c = *p;
    if (c == '(') {
      stack.push_back(Open{0, {0, 0, 0}, {0, 0, 0}});
      ++p;
    } else if (c == ',') {
      ++p;
    } else if (c == ')') {
      ++p;
      if (stack.empty()) return "unbalanced ')'";
      const Open o = stack.back();
      stack.pop_back();
      p = skip(p);
      while (!is_delim(*p)) ++p;  // inner label, ignored
      if (stack.empty()) {
        // the trifurcation (a trailing ":len" on it is accepted and ignored)
        double ignored;
        const char* q = branch_length(p, end, &ignored);
        if (!q) return "malformed branch length";
        p = skip(q);
        if (*p != ';') return "expected ';' after the top-level node";
        if (o.n != 3) return "top level must have exactly three subtrees";
        if (emitted != n_ops - 1) return "wrong number of tips for this family";
        children[2 * emitted] = o.id[1];
        children[2 * emitted + 1] = o.id[2];
        blen[2 * emitted] = o.len[1];
        blen[2 * emitted + 1] = o.len[2];
        ++emitted;
        *root_child = o.id[0];
        *root_blen = o.len[0];
        done = true;
      } else {
        if (o.n != 2) return "inner nodes must be binary";
        if (emitted >= n_ops - 1) return "wrong number of tips for this family";
        double len;
        p = branch_length(p, end, &len);
        if (!p) return "malformed branch length";
        children[2 * emitted] = o.id[0];
        children[2 * emitted + 1] = o.id[1];
        blen[2 * emitted] = o.len[0];
        blen[2 * emitted + 1] = o.len[1];
        Open& up = stack.back();
        if (up.n >= 3) return "node with more than three children";
        up.id[up.n] = T + emitted;
        up.len[up.n] = len;
        ++up.n;
        ++emitted;
      }
    } else if (c == '\0' || c == ';' || c == ':') {
      return "unexpected end of tree";
    } else {
      const char* b = p;
      uint32_t h = 2166136261u;
      for (; !is_delim(*p); ++p) h = (h ^ (unsigned char)*p) * 16777619u;
      if (stack.empty()) return "tip outside parentheses";
      const int32_t row = labels.find(b, (uint32_t)(p - b), h);
      if (row < 0) {
        detail.assign(b, p - b);
        return "tip label is not a sequence of this family";
      }
      if (seen[row]) {
        detail.assign(b, p - b);
        return "tip label appears twice";
      }
      seen[row] = 1;
      double len;
      p = branch_length(p, end, &len);
      if (!p) return "malformed branch length";
      Open& up = stack.back();
      if (up.n >= 3) return "node with more than three children";
      up.id[up.n] = row;
      up.len[up.n] = len;
      ++up.n;
    }
  }
  for (int i = 0; i < T; ++i)
    if (!seen[i]) return "tree is missing a sequence of this family";
  return nullptr;
}

}  // namespace
}  // namespace lhb

extern "C" int lhh_newick_to_ops(int32_t n_tips, const char* const* labels, int32_t n_trees,
                                 const char* const* newicks, int32_t n_threads, int32_t* children,
                                 double* blen, int32_t* root_child, double* root_blen) {
  using namespace lhb;
  if (n_tips < 3) return host_fail("need at least 3 tips (got %d)", n_tips);
  LabelMap map;
  map.init(n_tips);
  for (int i = 0; i < n_tips; ++i)
    if (!map.insert(labels[i], i)) return host_fail("duplicate label '%s'", labels[i]);
  const size_t per = (size_t)(n_tips - 2) * 2;
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  nt = std::max(1, std::min(nt, std::min(64, (n_trees + 63) / 64)));
  std::atomic<int> bad(-1);
  std::vector<std::string> msg(nt);
  auto work = [&](int tid) {
    std::vector<Open> stack;
    std::vector<uint8_t> seen;
    stack.reserve(256);
    for (int t = tid; t < n_trees; t += nt) {
      std::string detail;
      const char* err = flatten_one(newicks[t], map, n_tips, stack, seen, children + t * per, blen + t * per,
                                    root_child + t, root_blen + t, detail);
      if (err) {
        int expect = -1;
        if (bad.compare_exchange_strong(expect, t))
          msg[tid] = std::string(err) + (detail.empty() ? "" : " ('" + detail + "')");
        return;
      }
    }
  };
  if (nt == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int i = 0; i < nt; ++i) th.emplace_back(work, i);
    for (auto& t : th) t.join();
  }
  if (bad.load() >= 0) {
    std::string m;
    for (auto& s : msg)
      if (!s.empty()) m = s;
    return host_fail("tree %d: %s", bad.load(), m.c_str());
  }
  return 0;
}

// The inverse, for workloads and tests: the trees of a flat batch written as RevBayes writes them
// ("[&index=N]" comments, shortest round-trip branch lengths). Tree t goes to out + t * stride,
// NUL-terminated; flattening the strings again gives the same arrays when the batch's operations are in
// listed (post-)order. Returns 1 when `stride` is too small for some tree.
extern "C" int lhh_ops_to_newick(int32_t n_tips, const char* const* labels, int32_t n_trees, const int32_t* children,
                                 const double* blen, const int32_t* root_child, const double* root_blen,
                                 int32_t n_threads, char* out, int64_t stride) {
  using namespace lhb;
  const int T = n_tips, n_ops = T - 2;
  if (T < 3) return host_fail("need at least 3 tips (got %d)", T);
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  nt = std::max(1, std::min(nt, std::min(64, (n_trees + 15) / 16)));
  std::atomic<int> bad(-1);
  auto work = [&](int tid) {
    // iterative expansion with an explicit stack of (node, next child) frames
    struct Frame { int32_t node; int8_t next; };
    std::vector<Frame> stack;
    for (int t = tid; t < n_trees; t += nt) {
      const int32_t* ch = children + (size_t)t * n_ops * 2;
      const double* bl = blen + (size_t)t * n_ops * 2;
      char* const base = out + (size_t)t * stride;
      char* p = base;
      char* const end = base + stride - 64;  // room for one more token
      int counter = 0;
      bool ok = true;
      auto put_len = [&](double v) {
        p += std::snprintf(p, 32, "[&index=%d]:", ++counter);
        p = std::to_chars(p, p + 32, v).ptr;
      };
      // node -> text, children in listed order
      auto emit = [&](int32_t root) {
        stack.clear();
        stack.push_back({root, 0});
        while (!stack.empty() && ok) {
          Frame& f = stack.back();
          if (p >= end) { ok = false; break; }
          if (f.node < T) {
            const char* l = labels[f.node];
            const size_t n = std::strlen(l);
            if (p + n >= end) { ok = false; break; }
            std::memcpy(p, l, n);
            p += n;
            stack.pop_back();
            continue;
          }
          const int op = f.node - T;
          if (f.next == 0) {
            *p++ = '(';
            f.next = 1;
            stack.push_back({ch[2 * op], 0});
          } else if (f.next == 1) {
            put_len(bl[2 * op]);
            *p++ = ',';
            f.next = 2;
            stack.push_back({ch[2 * op + 1], 0});
          } else {
            put_len(bl[2 * op + 1]);
            *p++ = ')';
            stack.pop_back();
          }
        }
      };
      *p++ = '(';
      emit(root_child[t]);
      if (ok) { put_len(root_blen[t]); *p++ = ','; emit(ch[2 * (n_ops - 1)]); }
      if (ok) { put_len(bl[2 * (n_ops - 1)]); *p++ = ','; emit(ch[2 * (n_ops - 1) + 1]); }
      if (ok) {
        put_len(bl[2 * (n_ops - 1) + 1]);
        p += std::snprintf(p, 32, ")[&index=%d];", ++counter);
        *p = '\0';
      } else {
        bad.store(t);
        return;
      }
    }
  };
  if (nt == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int i = 0; i < nt; ++i) th.emplace_back(work, i);
    for (auto& t : th) t.join();
  }
  if (bad.load() >= 0) return host_fail("tree %d does not fit %lld bytes", bad.load(), (long long)stride);
  return 0;
}
