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

/* The inverse, for workloads and tests: the trees of a flat batch as RevBayes-style Newick text
 * ("[&index=N]" comments, shortest round-trip lengths); tree t is written at out + t * stride,
 * NUL-terminated. Fails when stride is too small.                                        */
int lhh_ops_to_newick(int32_t n_tips, const char* const* labels, int32_t n_trees, const int32_t* children,
                      const double* blen, const int32_t* root_child, const double* root_blen,
                      int32_t n_threads, char* out, int64_t stride);

/* Test hook: the decimal -> double conversion of the device Newick parser (csrc/decimal.hpp, Eisel-Lemire),
 * compiled for the host. n NUL-terminated strings laid end to end in buf; status 0 = converted (out is the
 * correctly rounded double, as atof in the reference's lexer), 1 = no number, 2 = left to the host parser;
 * consumed = characters read.                                                                    */
int lhh_parse_decimal_buf(const char* buf, int32_t n, double* out, int32_t* status, int32_t* consumed);

/* ---- linearham::PhyloHMM (linearham_b200/csrc/host/phylo_hmm.hpp), the host-side mirror of the
 * reference class (src/PhyloHMM.hpp:24-127, src/HMM.hpp:31-413), behind a C handle so that
 * non-C++ callers (the ctypes tests, bench.py) can drive it. Construction is pure host work
 * (YAML -> state space -> transitions -> xMSA) and needs no GPU; the scoring calls go through the
 * device ABI and fail without one.                                                            */
struct lh_family_desc;
typedef struct lhh_phylohmm lhh_phylohmm_t;

/* PhyloHMM(yaml_path, cluster_ind, hmm_param_dir, seed) - src/PhyloHMM.cpp:29-34. NULL on failure. */
lhh_phylohmm_t* lhh_phylohmm_new(const char* yaml_path, int32_t cluster_ind, const char* hmm_param_dir, int32_t seed);
void lhh_phylohmm_free(lhh_phylohmm_t* h);
int lhh_phylohmm_set_device(lhh_phylohmm_t* h, int32_t device);
/* Several GPUs of the box (before the first scoring call): the batched calls - score_samples, run_pipeline - cut
 * their trees into one shard per device (lh_group_t, include/linearham_b200.h); results do not depend on it. */
int lhh_phylohmm_set_devices(lhh_phylohmm_t* h, const int32_t* devices, int32_t n_devices);
/* Host threads this object may use (Newick ingest and traversal planning); 0 = all hardware threads.
 * A job with several ranks per node gives each rank its share. */
int lhh_phylohmm_set_host_threads(lhh_phylohmm_t* h, int32_t n_threads);
/* run_pipeline: 1 (default) = the states of every row's annotation are drawn on the device from the forward tables
 * left there (lh_sample_paths; one device only), 0 = by the host's SampleNaiveSequence; the output is the same.   */
int lhh_phylohmm_set_device_sampler(lhh_phylohmm_t* h, int32_t on);
/* The flat family description for lh_family_create; pointers live as long as the handle. */
int lhh_phylohmm_family_desc(lhh_phylohmm_t* h, struct lh_family_desc* out);
/* InitializePhyloParameters(newick_path, er, pi, alpha, num_rates) - src/PhyloHMM.cpp:350-361 */
int lhh_phylohmm_init_params(lhh_phylohmm_t* h, const char* newick_path, const double* er /*[6]*/,
                             const double* pi /*[4]*/, double alpha, int32_t num_rates);
int lhh_phylohmm_init_emission(lhh_phylohmm_t* h);             /* InitializePhyloEmission(), :366-383 */
int lhh_phylohmm_loglikelihood(lhh_phylohmm_t* h, double* out); /* LogLikelihood(), src/HMM.cpp:345-354 */
int lhh_phylohmm_sample_naive(lhh_phylohmm_t* h, char* out, int32_t cap); /* SampleNaiveSequence(), :358-431 */
/* RunPipeline(input_path, output_path, num_rates) - src/PhyloHMM.cpp:393-446 - scoring batch_size
 * RevBayes samples per device call.                                                            */
int lhh_phylohmm_run_pipeline(lhh_phylohmm_t* h, const char* input_path, const char* output_path,
                              int32_t num_rates, int32_t batch_size);
/* The loop body :419-430 for n samples at once: log-likelihoods only.                           */
int lhh_phylohmm_score_samples(lhh_phylohmm_t* h, int32_t n, const char* const* newicks, const double* er,
                               const double* pi, const double* alpha, int32_t num_rates, double* loglik);
/* The same with the strings' lengths (NULL: as above). Strings in pinned memory (lh_pin_host), one every constant
 * number of bytes, are copied to the device where they lie.                                      */
int lhh_phylohmm_score_samples_len(lhh_phylohmm_t* h, int32_t n, const char* const* newicks, const int64_t* lengths,
                                   const double* er, const double* pi, const double* alpha, int32_t num_rates,
                                   double* loglik);
/* The device handle (include/linearham_b200.h) behind the object, created on first use: for
 * lh_get_timings / lh_set_option / lh_loglik_dev after a scoring call. Owned by the object.       */
struct lh_family;
int lhh_phylohmm_device_family(lhh_phylohmm_t* h, struct lh_family** out);
/* pll_utree_export_newick of a Newick string (libpll utree.c:148-270), as written to the output TSV. */
int lhh_export_newick(const char* newick, char* out, int32_t cap);
/* Named read-outs of the object's members (accessor names of src/HMM.hpp / src/PhyloHMM.hpp
 * without the trailing underscore), column-major; out = NULL queries the shape.                 */
int lhh_phylohmm_get_f64(lhh_phylohmm_t* h, const char* name, double* out, int32_t* rows, int32_t* cols);
int lhh_phylohmm_get_i32(lhh_phylohmm_t* h, const char* name, int32_t* out, int32_t* rows, int32_t* cols);
int lhh_phylohmm_get_str(lhh_phylohmm_t* h, const char* name, char* out, int32_t cap);

#ifdef __cplusplus
}
#endif
#endif /* LINEARHAM_B200_HOST_H_ */
