// k_plan: the traversal planner on the device. Same algorithm and bit-identical programs as the host
// planner (plan.cpp; replaces pll_utree_traverse + pll_utree_create_operations, utree.c:284-442), so
// that lh_score_batch ships the caller's operation lists as they are (24 B per operation instead of
// the 32 B of a planned step) and the host does no per-tree work beyond a memcpy into pinned memory.
//
// One warp per tree, three tables in shared memory (children, `up`, `down` of the inner nodes: 12 B each):
//   1. all lanes: copy the operation list in, check branch lengths, count references (every tip and
//      every inner node is consumed exactly once), check that children precede parents;
//   2. lane 0: the bottom-up sweep (need, size per node) - a dependent chain through `up`;
//   3. lane 0: the top-down sweep hands every node its block start, parking slot, level and parity;
//   4. all lanes: every operation builds its own step from the tables and stores it at its program
//      position, with its two branch lengths (coalesced reads, scattered 16-byte writes).
// A malformed tree (or a negative branch length) is reported through `flags` and gets a harmless
// program; the host turns the flag into the error of the call.
#include <cuda_runtime.h>

#include <climits>
#include <cstdint>
#include <map>
#include <mutex>

#include "kernels.hpp"

namespace lhb {

namespace {

__device__ __forceinline__ uint32_t pack_down(uint32_t start, int want, uint32_t level, uint32_t par) {
  return start << 17 | (uint32_t)(want & 0xff) | level << 8 | par << 16;
}

// Shared memory through 32-bit shared-space addresses (through generic pointers every access in the
// lane-0 sweeps rebuilt the CTA's shared window base with an S2R on the dependent chain).
__device__ __forceinline__ uint32_t lds(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v));
}

__global__ void __launch_bounds__(32) k_plan(PlanDev p) {
  extern __shared__ uint32_t sm[];
  const int T = p.T, n_ops = T - 2, n_nodes = T + n_ops, lane = threadIdx.x;
  // tables of the inner nodes only (a tip's `up` word is 0 and nobody reads its `down` word);
  // node ids fit 16 bits (plan_on_device_ok), so an operation's two children are one word
  const uint32_t s_up = (uint32_t)__cvta_generic_to_shared(sm);  // [n_ops] need | size << 8
  const uint32_t s_dn = s_up + 4u * n_ops;                       // [n_ops] reference counts, then the down word
  const uint32_t s_ch = s_dn + 4u * n_ops;                       // [n_ops] child a | child b << 16
  uint32_t* dn_gen = sm + n_ops;                                 // (generic view, for the atomics)
  uint32_t* tip_seen = sm + 3 * n_ops;                           // [(T+31)/32] one bit per tip: referenced
  const int t = p.tree0 + blockIdx.x;
  const int32_t* __restrict__ gch = p.children + (size_t)t * 2 * n_ops;
  const double* __restrict__ gbl = p.blen + (size_t)t * 2 * n_ops;
  int4* __restrict__ out = p.plan + (size_t)t * (T - 1);
  double* __restrict__ pb = p.pblen + (size_t)t * (2 * T - 3);
  const int32_t rc = p.root_child[t];
  const double rbl = p.root_blen[t];

  // 1. load + checks: children precede parents, every inner node is consumed exactly once
  bool neg = !(rbl >= 0.0), bad = false;
  for (int i = lane; i < n_ops; i += 32) sts(s_dn + 4u * i, 0u);
  for (int i = lane; i < (T + 31) / 32; i += 32) tip_seen[i] = 0u;
  __syncwarp();
  // a tip referenced twice shows as a bit already set; with every inner node consumed once the 2T-3
  // references leave exactly T for the T tips, so "no tip twice" is "every tip once"
  auto see_tip = [&](int id) { return (atomicOr(&tip_seen[id >> 5], 1u << (id & 31)) >> (id & 31)) & 1u; };
#pragma unroll 4
  for (int op = lane; op < n_ops; op += 32) {
    const int2 c = *reinterpret_cast<const int2*>(gch + 2 * op);
    const double2 l = *reinterpret_cast<const double2*>(gbl + 2 * op);
    neg |= !(l.x >= 0.0) | !(l.y >= 0.0);
    if ((uint32_t)c.x >= (uint32_t)(T + op) || (uint32_t)c.y >= (uint32_t)(T + op)) {
      bad = true;
    } else {
      sts(s_ch + 4u * op, (uint32_t)c.x | (uint32_t)c.y << 16);
      if (c.x >= T) atomicAdd(&dn_gen[c.x - T], 1u); else bad |= see_tip(c.x) != 0u;
      if (c.y >= T) atomicAdd(&dn_gen[c.y - T], 1u); else bad |= see_tip(c.y) != 0u;
    }
  }
  if (rc < 0 || rc >= n_nodes - 1) bad = true;
  else if (lane == 0) {
    if (rc >= T) atomicAdd(&dn_gen[rc - T], 1u); else bad |= see_tip(rc) != 0u;
  }
  __syncwarp();
  for (int i = lane; i < n_ops - 1; i += 32) bad |= lds(s_dn + 4u * i) != 1u;
  if (lane == 0) bad |= lds(s_dn + 4u * (n_ops - 1)) != 0u;
  neg = __any_sync(0xffffffffu, neg);
  bad = __any_sync(0xffffffffu, bad);
  if (neg || bad) {
    if (lane == 0) atomicMin(p.flags + (neg ? 0 : 1), t);
    for (int i = lane; i < n_ops; i += 32) out[i] = make_int4(1, 2, -1, 0);
    if (lane == 0) out[n_ops] = make_int4(-1, 1, -1, 0);
    for (int i = lane; i < 2 * n_ops + 1; i += 32) pb[i] = 0.0;
    return;
  }
  auto up_of = [&](int id) -> uint32_t { return id >= T ? lds(s_up + 4u * (id - T)) : 0u; };

  // 2. bottom-up: need (live CLVs, 0 for a tip) and size (inner nodes) of every subtree.
  // The children of the next operation are fetched while this one is computed: the chain through
  // `up` is then one shared-memory load per operation.
  if (lane == 0) {
    uint32_t c_next = lds(s_ch);
    for (int op = 0; op < n_ops; ++op) {
      const uint32_t c = c_next;
      if (op + 1 < n_ops) c_next = lds(s_ch + 4u * (op + 1));
      const uint32_t ua = up_of(c & 0xffff), ub = up_of(c >> 16);
      const uint32_t na = ua & 0xff, nb = ub & 0xff;
      const uint32_t n = na == nb ? na + 1 : max(na, nb);
      sts(s_up + 4u * op, n | ((ua >> 8) + (ub >> 8) + 1) << 8);
    }
  }
  __syncwarp();

  // 3. top-down: block start, parking slot, level, parity for every inner node
  const int root_op = n_ops - 1;
  int4 final_step = make_int4(-1, rc, -1, 0);
  int deepest = 0;
  if (rc < T) {
    if (lane == 0) sts(s_dn + 4u * root_op, pack_down(0, -1, 0, 0));
  } else {
    const uint32_t u_rc = lds(s_up + 4u * (rc - T)), u_root = lds(s_up + 4u * root_op);
    if ((u_rc & 0xff) >= (u_root & 0xff)) {  // root_child first, parked in slot 0
      final_step = make_int4(-1, -2, -1, 0);
      if (lane == 0) {
        sts(s_dn + 4u * (rc - T), pack_down(0, 0, 0, 0));
        sts(s_dn + 4u * root_op, pack_down(u_rc >> 8, -1, 1, 0));
      }
    } else {                                 // the virtual root first, parked in slot 0
      final_step = make_int4(-2, -1, -1, 0);
      if (lane == 0) {
        sts(s_dn + 4u * root_op, pack_down(0, 0, 0, 0));
        sts(s_dn + 4u * (rc - T), pack_down(u_root >> 8, -1, 1, 0));
      }
    }
    deepest = 1;
  }
  if (lane == 0) {
    uint32_t c_next = lds(s_ch + 4u * (n_ops - 1));
    for (int op = n_ops - 1; op >= 0; --op) {
      const uint32_t c = c_next;
      if (op > 0) c_next = lds(s_ch + 4u * (op - 1));
      const int a = c & 0xffff, b = c >> 16;
      const uint32_t d = lds(s_dn + 4u * op);
      const uint32_t ua = up_of(a), ub = up_of(b);
      const uint32_t level = (d >> 8) & 0xff, begin = d >> 17;
      const uint32_t both = (uint32_t)(a >= T) & (uint32_t)(b >= T);
      const uint32_t a_first = (ua & 0xff) >= (ub & 0xff);
      const uint32_t a_parks = both & a_first, b_parks = both & (a_first ^ 1u);
      const uint32_t base = (d & 0x1ff00u) ^ 0x10000u;  // level and flipped parity
      const uint32_t lo_a = base + (b_parks << 8) + (a_parks ? level : 0xffu);
      const uint32_t lo_b = base + (a_parks << 8) + (b_parks ? level : 0xffu);
      const uint32_t st_a = begin + (a_first ? 0u : ub >> 8), st_b = begin + (a_first ? ua >> 8 : 0u);
      if (a >= T) sts(s_dn + 4u * (a - T), st_a << 17 | lo_a);
      if (b >= T) sts(s_dn + 4u * (b - T), st_b << 17 | lo_b);
    }
  }
  __syncwarp();

  // 4. every operation writes its step at its program position
#pragma unroll 4
  for (int op = lane; op < n_ops; op += 32) {
    const uint32_t c = lds(s_ch + 4u * op);
    const int a = c & 0xffff, b = c >> 16;
    const double2 l = *reinterpret_cast<const double2*>(gbl + 2 * op);
    const uint32_t ua = up_of(a), ub = up_of(b), d = lds(s_dn + 4u * op);
    const uint32_t level = (d >> 8) & 0xff, begin = d >> 17;
    const uint32_t pos = begin + (lds(s_up + 4u * op) >> 8) - 1;
    const uint32_t ia = a >= T, ib = b >= T, both = ia & ib;
    const uint32_t a_first = (ua & 0xff) >= (ub & 0xff);
    const uint32_t a_parks = both & a_first, b_parks = both & (a_first ^ 1u);
    const int32_t src_a = ia ? -1 - (int32_t)(a_parks * (level + 1)) : a;
    const int32_t src_b = ib ? -1 - (int32_t)(b_parks * (level + 1)) : b;
    out[pos] = make_int4(src_a, src_b, (int32_t)(int8_t)(d & 0xff), (int32_t)((d >> 16) & 1));
    pb[2 * pos] = l.x;
    pb[2 * pos + 1] = l.y;
    deepest = max(deepest, (int)(level + both));
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) deepest = max(deepest, __shfl_xor_sync(0xffffffffu, deepest, o));
  if (lane == 0) {
    out[n_ops] = final_step;
    pb[2 * n_ops] = rbl;
    atomicMax(p.flags + 2, deepest);
  }
}

}  // namespace

void ensure_dynamic_smem(const void* kernel, size_t bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, size_t> limit;  // the attribute is per device
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> g(mu);
  size_t& cur = limit[{dev, kernel}];
  if (bytes > cur) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess)
      cur = bytes;
    else
      cudaGetLastError();  // the launch that needs it will report
  }
}

size_t plan_smem_bytes(int n_tips) {
  return (size_t)(n_tips - 2) * (2 * sizeof(uint32_t) + 2 * sizeof(uint16_t)) + (size_t)(n_tips + 31) / 32 * sizeof(uint32_t);
}

// The device planner packs a block start into 15 bits and keeps three tables in shared memory.
bool plan_on_device_ok(int n_tips) { return n_tips - 2 < (1 << 15) && plan_smem_bytes(n_tips) <= (size_t)200 * 1024; }

void launch_plan(const PlanDev& p, int ntrees, cudaStream_t s) {
  if (ntrees <= 0) return;
  const size_t smem = plan_smem_bytes(p.T);
  ensure_dynamic_smem(reinterpret_cast<const void*>(&k_plan), smem);
  k_plan<<<ntrees, 32, smem, s>>>(p);
}

}  // namespace lhb
