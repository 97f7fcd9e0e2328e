/* linearham_b200 - host-side C ABI (no CUDA): tree ingest, Gamma categories, traversal
 * planning and the C wrapper over the host PhyloHMM class. These are the pieces of the
 * reference's per-sample host work that stay on the CPU:
 *   Newick -> operations      libpll parse_utree.y:179-445, utree.c:284-442,
 *                             src/PhyloHMM.cpp:419-422
 *   discrete-Gamma rates      libpll gamma.c:220-292 (src/PhyloHMM.cpp:425-426)
 * plus test hooks for the traversal planner (linearham_b200/csrc/plan.hpp).
 * All functions return 0 on success; lhh_last_error() describes the last failure.   */
#ifndef LINEARHAM_B200_HOST_H_
#define LINEARHAM_B200_HOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* lhh_last_error(void);

/* Traversal planner (test hook). steps: [B][T-1][4] int32 (a, b, dst, 0);
 * pblen: [B][2T-3]; max_slots: stack depth the batch needs.                          */
int lhh_plan_trees(int32_t n_tips, int32_t n_trees, const int32_t* children, const double* blen,
                   const int32_t* root_child, const double* root_blen, int32_t* steps,
                   double* pblen, int32_t* max_slots);

#ifdef __cplusplus
}
#endif
#endif /* LINEARHAM_B200_HOST_H_ */
