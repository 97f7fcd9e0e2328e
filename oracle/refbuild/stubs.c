/* TEST INFRASTRUCTURE ONLY (oracle/). Symbols libptpll references from
 * optimiser / Laplace-marginal code paths that linearham never calls
 * (lib/libptpll/src/pll_partition.cpp:261-307,408-752). They need GSL and the
 * pll-modules optimisers, which cannot be built here; calling one aborts. */
#include <stdio.h>
#include <stdlib.h>
#define LH_STUB(name) \
  void name(void) { fprintf(stderr, "oracle stub reached: %s\n", #name); abort(); }
LH_STUB(lesplace_gamma)
LH_STUB(pllmod_algo_opt_alpha)
LH_STUB(pllmod_algo_opt_frequencies)
LH_STUB(pllmod_algo_opt_subst_rates)
LH_STUB(pllmod_opt_optimize_branch_lengths_iterative)
LH_STUB(pllmod_opt_optimize_branch_lengths_local)
