// Host-side traversal planner: turns one tree's operation list (reference
// post-order, libpll utree.c:284-329) into the register/stack program the
// pruning kernel executes.
//
// The reference materialises every inner CLV in its own buffer
// (pll_partition_create, pll.c:525-542). Here an inner CLV only ever lives in
// (a) the registers of the thread that owns the site ("REG", the result of the
// previous step) or (b) a small per-site stack of slots. Children of a node are
// evaluated larger-subtree-first (Sethi-Ullman / Strahler order), which bounds
// the stack depth by floor(log2(T)) and leaves every result that is consumed by
// the very next step in registers. Re-ordering is value-preserving: the parent
// is a product of two child terms (core_partials.c:620-628) and IEEE
// multiplication and integer scaler addition commute.
#pragma once
#include <cstdint>
#include <vector>

namespace lhb {

// Child source encoding in a step.
//   >= 0 : tip row (row of tip_codes)
//   -1   : REG (result of the previous step)
//   <= -2: stack slot (-2 - src)
constexpr int32_t kSrcReg = -1;
inline int32_t src_slot(int slot) { return -2 - slot; }

struct Step {          // 16 bytes, loaded as one int4 by the kernel
  int32_t a, b;        // sources of the two children (for the final step: parent, child)
  int32_t dst;         // slot to copy the result into, or -1
  int32_t pad;         // node steps: parity of the node's depth below the root edge
};

// Plans one tree.
//   children/blen : [T-2][2] as in lh_tree_batch_t
//   steps         : [T-1] out; steps[0..T-3] are node updates, steps[T-2] is the
//                   edge evaluation (a = virtual-root CLV source, b = root_child source)
//   pblen         : [2(T-2)+1] out; branch length of step s child c at 2s+c, root edge last
//   scratch       : reusable work buffer
// Returns the number of stack slots the program needs, or -1 for a malformed tree.
struct Step;
struct PlanScratch {
  std::vector<uint32_t> up;
  std::vector<int32_t> order;
  std::vector<Step> by_op;
  std::vector<uint64_t> down;
  std::vector<uint8_t> uses;
};
int plan_tree(int n_tips, const int32_t* children, const double* blen, int32_t root_child,
              double root_blen, Step* steps, double* pblen, PlanScratch& scratch);

}  // namespace lhb
