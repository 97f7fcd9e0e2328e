#include "plan.hpp"

#include <emmintrin.h>

#include <algorithm>
#include <cstring>

namespace lhb {

int plan_tree(int T, const int32_t* children, const double* blen, int32_t root_child,
              double root_blen, Step* steps, double* pblen, PlanScratch& sc) {
  const int n_ops = T - 2, n_nodes = T + n_ops;
  if (T < 3 || n_ops >= (1 << 24)) return -1;
  // Which child of a node is a tip is a coin toss, so every per-node decision below is written
  // as arithmetic on arrays indexed by node id (tips included) rather than as a branch, and
  // what is known about a node is packed into one word per direction to keep the tables in L1.

  // up[id]: facts a node's parent reads, filled in operation (= post-) order.
  //   need  bits 0..7   number of simultaneously live CLVs (REG included) to evaluate the node; 0 for
  //                     a tip, which makes the three cases one rule: equal needs -> one more, else the larger
  //   size  bits 8..31  inner nodes in the subtree of id (0 for a tip)
  // uses[id] counts references: every tip and every inner node except the virtual root must be consumed
  // exactly once (root_child by the root edge).
  sc.up.assign(n_nodes, 0);
  sc.uses.assign(n_nodes, 0);
  uint32_t* up = sc.up.data();
  uint8_t* uses = sc.uses.data();
  for (int op = 0; op < n_ops; ++op) {
    const int32_t a = children[2 * op], b = children[2 * op + 1];
    if ((uint32_t)a >= (uint32_t)(T + op) || (uint32_t)b >= (uint32_t)(T + op)) return -1;  // not a post-order reference
    uses[a] += uses[a] < 2;  // saturating: "more than once" is all that matters
    uses[b] += uses[b] < 2;
    const uint32_t ua = up[a], ub = up[b];
    const uint32_t na = ua & 0xff, nb = ub & 0xff;
    const uint32_t n = na == nb ? na + 1 : std::max(na, nb);
    up[T + op] = n | ((ua >> 8) + (ub >> 8) + 1) << 8;
  }
  if (root_child < 0 || root_child >= n_nodes - 1) return -1;
  uses[root_child] += uses[root_child] < 2;
  int wrong = uses[n_nodes - 1];
  for (int id = 0; id < n_nodes - 1; ++id) wrong |= uses[id] ^ 1;  // every tip and inner node consumed exactly once
  if (wrong) return -1;
  auto need = [&](int32_t id) { return (int)(up[id] & 0xff); };
  auto size = [&](int32_t id) { return (int)(up[id] >> 8); };

  // Execution order = post-order with the child that needs more live CLVs first. The operations
  // come in post-order, so a parent always has a higher index than its children and one sweep
  // from the last operation down hands every node what it needs to place itself - no traversal
  // stack, no dependence between neighbouring iterations except through down[id]:
  //   start bits 32..63 index of the first step of its subtree's block in the program; its own step is
  //         the block's last (start + size - 1). The child evaluated first starts where its
  //         parent's block starts, the other one right after that child's block.
  //   want  bits 0..7   slot its result is also copied to (0xff = -1: stays in REG only)
  //   level bits 8..15  stack depth when its subtree starts = the slot its first-evaluated child
  //         parks in. A parked result occupies its slot only from the moment it is produced, so
  //         the first child runs at `level`, the second at `level + 1`.
  //   par   bit 16      parity of the node's distance from the root edge
  sc.down.resize(n_nodes);
  uint64_t* down = sc.down.data();
  auto pack = [](int start, int want, int level, int par) {
    return (uint64_t)(uint32_t)start << 32 | (uint32_t)(want & 0xff) | (uint32_t)level << 8 | (uint32_t)par << 16;
  };
  int max_depth = 0;
  const int32_t root_node = n_nodes - 1;
  Step final_step{kSrcReg, root_child, -1, 0};
  if (root_child < T) {
    down[root_node] = pack(0, -1, 0, 0);
  } else if (need(root_child) >= need(root_node)) {  // root_child first, parked in slot 0
    final_step = {kSrcReg, src_slot(0), -1, 0};
    down[root_child] = pack(0, 0, 0, 0);
    down[root_node] = pack(size(root_child), -1, 1, 0);
    max_depth = 1;
  } else {                                           // the virtual root first, parked in slot 0
    final_step = {src_slot(0), kSrcReg, -1, 0};
    down[root_node] = pack(0, 0, 0, 0);
    down[root_child] = pack(size(root_node), -1, 1, 0);
    max_depth = 1;
  }
  // The sweep only writes small per-thread tables (the step of every operation, in operation order,
  // and which operation runs at which program position); the program itself is then written front to
  // back. Scattering the steps straight into the caller's (cold, pinned) buffer would make every one
  // of them a cache miss that the in-order store queue has to wait for.
  sc.by_op.resize(n_ops);
  sc.order.resize(n_ops);
  {
    const uint32_t* __restrict upr = up;
    uint64_t* __restrict dnr = down;
    const int32_t* __restrict ch = children;
    Step* __restrict by_op = sc.by_op.data();
    int32_t* __restrict order = sc.order.data();
    int deepest = max_depth;
    for (int op = n_ops - 1; op >= 0; --op) {
      const int32_t a = ch[2 * op], b = ch[2 * op + 1];
      const uint32_t ua = upr[a], ub = upr[b];
      const uint64_t dn = dnr[T + op];
      const uint32_t lo = (uint32_t)dn;                     // want | level << 8 | par << 16
      const uint32_t level = (lo >> 8) & 0xff, begin = (uint32_t)(dn >> 32);
      const uint32_t pos = begin + (upr[T + op] >> 8) - 1;
      const uint32_t ia = a >= T, ib = b >= T, both = ia & ib;
      const uint32_t a_first = (ua & 0xff) >= (ub & 0xff);  // a tip needs 0: the inner child of a mixed pair is "first"
      const uint32_t a_parks = both & a_first, b_parks = both & (a_first ^ 1);
      // sources: tip id | REG (-1) | the slot the first child parked in (-2 - level): REG - parks * (level + 1)
      const int32_t src_a = ia ? -1 - (int32_t)(a_parks * (level + 1)) : a;
      const int32_t src_b = ib ? -1 - (int32_t)(b_parks * (level + 1)) : b;
      by_op[op] = Step{src_a, src_b, (int32_t)(int8_t)lo, (int32_t)((lo >> 16) & 1)};
      order[pos] = op;
      deepest = std::max(deepest, (int)(level + both));
      // children: level and flipped parity from the parent, want = level if parked else 0xff;
      // the second-evaluated child runs one level up and starts after the first one's block
      const uint32_t base = (lo & 0x1ff00u) ^ 0x10000u;
      const uint32_t lo_a = base + (b_parks << 8) + (a_parks ? level : 0xffu);
      const uint32_t lo_b = base + (a_parks << 8) + (b_parks ? level : 0xffu);
      const uint32_t st_a = begin + (a_first ? 0u : ub >> 8), st_b = begin + (a_first ? ua >> 8 : 0u);
      dnr[a] = (uint64_t)st_a << 32 | lo_a;   // (tips are written too, harmlessly)
      dnr[b] = (uint64_t)st_b << 32 | lo_b;
    }
    max_depth = deepest;
    const double* __restrict bl = blen;
    Step* __restrict out = steps;
    double* __restrict pb = pblen;
    // The destination is normally a pinned staging buffer the GPU reads next: non-temporal stores keep
    // the lines out of this core's cache (a DMA read of dirty lines runs at a fraction of the link rate).
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (reinterpret_cast<uintptr_t>(pb) & 7) == 0) {
      for (int pos = 0; pos < n_ops; ++pos) {
        const int op = order[pos];
        _mm_stream_si128(reinterpret_cast<__m128i*>(out + pos),
                         _mm_loadu_si128(reinterpret_cast<const __m128i*>(by_op + op)));
        long long l0, l1;
        std::memcpy(&l0, bl + 2 * op, 8);
        std::memcpy(&l1, bl + 2 * op + 1, 8);
        _mm_stream_si64(reinterpret_cast<long long*>(pb + 2 * pos), l0);
        _mm_stream_si64(reinterpret_cast<long long*>(pb + 2 * pos + 1), l1);
      }
      _mm_sfence();
    } else {
      for (int pos = 0; pos < n_ops; ++pos) {
        const int op = order[pos];
        out[pos] = by_op[op];
        pb[2 * pos] = bl[2 * op];
        pb[2 * pos + 1] = bl[2 * op + 1];
      }
    }
  }
  if (max_depth > 127) return -1;
  steps[n_ops] = final_step;
  pblen[2 * n_ops] = root_blen;
  return max_depth;
}

}  // namespace lhb
