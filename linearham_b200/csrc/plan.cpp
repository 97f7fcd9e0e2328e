#include "plan.hpp"

#include <algorithm>

namespace lhb {

namespace {
struct Frame {
  int32_t node;       // inner node id (>= T)
  int32_t want_slot;  // slot the result must also be copied to, or -1
  int32_t held_slot;  // slot holding the first-evaluated child (phase 1), or -1
  int32_t held_is_a;  // 1 if the held child is child a
  int32_t phase;      // 0 = expand, 1 = emit the node's own step, 2 = occupy one slot
  int32_t depth;      // edges between the node and the root edge
};
}  // namespace

int plan_tree(int T, const int32_t* children, const double* blen, int32_t root_child,
              double root_blen, Step* steps, double* pblen, PlanScratch& sc) {
  const int n_ops = T - 2;
  if (T < 3) return -1;
  auto is_inner = [T](int32_t id) { return id >= T; };

  // need[op] = number of simultaneously live CLVs (REG included) to evaluate the node
  sc.need.assign(n_ops, 0);
  std::vector<int32_t>& need = sc.need;
  // use count check: every inner node except the virtual root and root_child must be
  // consumed exactly once; sc.stack doubles as the counter array first.
  sc.stack.assign(n_ops, 0);
  for (int op = 0; op < n_ops; ++op) {
    int na = 0, nb = 0;
    for (int c = 0; c < 2; ++c) {
      int32_t id = children[2 * op + c];
      if (id < 0 || id >= T + op) return -1;  // not a valid post-order reference
      if (is_inner(id)) {
        if (++sc.stack[id - T] > 1) return -1;
        (c ? nb : na) = need[id - T];
      }
    }
    const bool ia = is_inner(children[2 * op]), ib = is_inner(children[2 * op + 1]);
    if (ia && ib)
      need[op] = (na == nb) ? na + 1 : std::max(na, nb);
    else if (ia || ib)
      need[op] = std::max(na, nb);
    else
      need[op] = 1;
  }
  if (root_child < 0 || root_child >= T + n_ops - 1) return -1;
  if (is_inner(root_child) && ++sc.stack[root_child - T] > 1) return -1;
  for (int op = 0; op + 1 < n_ops; ++op)
    if (sc.stack[op] != 1) return -1;
  if (sc.stack[n_ops - 1] != 0) return -1;

  int depth = 0, max_depth = 0, n_steps = 0;
  std::vector<Frame> fr;
  fr.reserve(64);
  // A parked result occupies its slot only from the moment it is produced: the slot index
  // is the stack depth when the producing subtree starts (its evaluation returns the depth
  // to that level), and the depth is bumped by a marker frame once the subtree is done.
  const Frame occupy{0, 0, 0, 0, 2, 0};

  const int32_t root_node = T + n_ops - 1;
  Step final_step{kSrcReg, root_child, -1, 0};
  // Frames are pushed in reverse order of execution.
  if (!is_inner(root_child)) {
    fr.push_back({root_node, -1, -1, 0, 0, 0});
  } else if (need[root_child - T] >= need[n_ops - 1]) {
    final_step = {kSrcReg, src_slot(0), -1, 0};
    fr.push_back({root_node, -1, -1, 0, 0, 0});
    fr.push_back(occupy);
    fr.push_back({root_child, 0, -1, 0, 0, 0});
  } else {
    final_step = {src_slot(0), kSrcReg, -1, 0};
    fr.push_back({root_child, -1, -1, 0, 0, 0});
    fr.push_back(occupy);
    fr.push_back({root_node, 0, -1, 0, 0, 0});
  }

  while (!fr.empty()) {
    Frame f = fr.back();
    fr.pop_back();
    if (f.phase == 2) {
      ++depth;
      max_depth = std::max(max_depth, depth);
      continue;
    }
    const int op = f.node - T;
    const int32_t a = children[2 * op], b = children[2 * op + 1];
    const bool ia = is_inner(a), ib = is_inner(b);
    if (f.phase == 0) {
      if (ia && ib) {
        const bool a_first = need[a - T] >= need[b - T];
        const int s = depth;
        fr.push_back({f.node, f.want_slot, s, a_first ? 1 : 0, 1, f.depth});
        fr.push_back({a_first ? b : a, -1, -1, 0, 0, f.depth + 1});  // second: ends in REG
        fr.push_back(occupy);
        fr.push_back({a_first ? a : b, s, -1, 0, 0, f.depth + 1});   // first: parked in slot s
      } else {
        fr.push_back({f.node, f.want_slot, -1, 0, 1, f.depth});
        if (ia) fr.push_back({a, -1, -1, 0, 0, f.depth + 1});
        if (ib) fr.push_back({b, -1, -1, 0, 0, f.depth + 1});
      }
      continue;
    }
    Step st;
    st.pad = f.depth & 1;  // parity of the node's depth: picks the CLV arena end (k_repeats)
    st.dst = f.want_slot;
    if (ia && ib) {
      st.a = f.held_is_a ? src_slot(f.held_slot) : kSrcReg;
      st.b = f.held_is_a ? kSrcReg : src_slot(f.held_slot);
      depth--;  // the parked child is consumed here
    } else {
      st.a = ia ? kSrcReg : a;
      st.b = ib ? kSrcReg : b;
    }
    pblen[2 * n_steps] = blen[2 * op];
    pblen[2 * n_steps + 1] = blen[2 * op + 1];
    steps[n_steps++] = st;
  }
  if (n_steps != n_ops) return -1;
  steps[n_ops] = final_step;
  pblen[2 * n_ops] = root_blen;
  return max_depth;
}

}  // namespace lhb
