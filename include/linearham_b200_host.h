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

/* Discrete-Gamma mean category rates; replaces pll_compute_gamma_cats(alpha, n, out,
 * PLL_GAMMA_RATES_MEAN) (libpll gamma.c:220-292; src/PhyloHMM.cpp:360,425-426). Same series,
 * tolerances and operation order, hence the same doubles. alpha < 0.02 is an error
 * (gamma.c:238). out: [n_rates]; the batch form fills out[n][n_rates].                */
int lhh_gamma_cats(double alpha, int32_t n_rates, double* out);
int lhh_gamma_cats_batch(int32_t n, const double* alpha, int32_t n_rates, double* out);

/* Newick strings -> flat operation arrays (the children/blen/root_child/root_blen members of
 * lh_tree_batch_t). Replaces regex_replace + pll_utree_parse_newick_string +
 * set_missing_branch_length + the post-order traversal (src/PhyloHMM.cpp:419-422,
 * parse_utree.y:179-246, pll_util.cpp:46-62, utree.c:284-442). labels[n_tips] are the xMSA
 * row labels (row 0 = "naive"). n_threads <= 0: one per hardware thread.             */
int lhh_newick_to_ops(int32_t n_tips, const char* const* labels, int32_t n_trees,
                      const char* const* newicks, int32_t n_threads, int32_t* children,
                      double* blen, int32_t* root_child, double* root_blen);

#ifdef __cplusplus
}
#endif
#endif /* LINEARHAM_B200_HOST_H_ */
