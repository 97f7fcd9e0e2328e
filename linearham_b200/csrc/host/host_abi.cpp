// Host-side C ABI (include/linearham_b200_host.h). No CUDA in this file.
#include "../../../include/linearham_b200_host.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../plan.hpp"
#include "host_error.hpp"

namespace lhb {
thread_local std::string g_host_err;
int host_fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_host_err = buf;
  return 1;
}
}  // namespace lhb

extern "C" {

const char* lhh_last_error(void) { return lhb::g_host_err.c_str(); }

int lhh_plan_trees(int32_t T, int32_t B, const int32_t* children, const double* blen,
                   const int32_t* root_child, const double* root_blen, int32_t* steps,
                   double* pblen, int32_t* max_slots) {
  static_assert(sizeof(lhb::Step) == 16, "Step must be an int4");
  lhb::PlanScratch sc;
  int mx = 0;
  for (int t = 0; t < B; ++t) {
    int s = lhb::plan_tree(T, children + (size_t)t * (T - 2) * 2, blen + (size_t)t * (T - 2) * 2,
                           root_child[t], root_blen[t],
                           reinterpret_cast<lhb::Step*>(steps) + (size_t)t * (T - 1),
                           pblen + (size_t)t * (2 * T - 3), sc);
    if (s < 0) return lhb::host_fail("tree %d is not a valid post-order operation list", t);
    if (s > mx) mx = s;
  }
  *max_slots = mx;
  return 0;
}

}  // extern "C"
