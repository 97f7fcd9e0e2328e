/* linearham_b200 - device C ABI for the phylo-HMM likelihood hot path.
 *
 * This is the drop-in boundary. The reference (matsengrp/linearham) has no FFI
 * for this path: its boundary is the C++ call sequence inside the RunPipeline
 * loop body,  src/PhyloHMM.cpp:419-430
 *
 *     InitializePhyloEmission();          // src/PhyloHMM.cpp:366-383
 *     lh_loglikelihood_ = LogLikelihood();// src/HMM.cpp:345-354
 *
 * executed once per (tree, er, pi, alpha) sample on a fixed clonal family.
 * The functions below replace that pair for B samples at a time; the host-side
 * PhyloHMM class of this repo (linearham_b200/csrc/host) calls nothing else.
 *
 * Conventions
 *   - plain C, host pointers unless a name ends in _dev; no torch / Eigen types
 *   - every call returns 0 on success, non-zero on failure; lh_last_error()
 *     then returns a thread-local message (the reference throws C++ exceptions
 *     caught in main(), src/linearham.cpp:447-454)
 *   - a handle is single-threaded; different handles may be used concurrently
 *   - there is NO CPU fallback: without a CUDA device lh_family_create fails
 *   - FP64 end to end; bases are 0..4 = A,C,G,T,N (src/HMM.cpp:50)
 */
#ifndef LINEARHAM_B200_H_
#define LINEARHAM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LH_MAX_RATES 8

typedef struct lh_family lh_family_t;

/* Per-family constants; replaces what PhyloHMM::InitializeXmsaStructs
 * (src/PhyloHMM.cpp:45-89), HMM::InitializeStateSpace/InitializeTransition
 * (src/HMM.cpp:94-246) and pt::pll::Partition::SetTipStates
 * (lib/libptpll/src/pll_partition.cpp:153-194) leave in host objects.
 * Matrices are dense column-major exactly as the reference's Eigen::MatrixXd /
 * MatrixXi members; the library builds its own sparse device forms.           */
typedef struct lh_family_desc {
  int32_t n_tips;            /* T = sequences + 1; row 0 is "naive"           */
  int32_t n_sites;           /* S_x = xMSA columns                            */
  const uint8_t* tip_codes;  /* [T][S_x] row-major, xmsa_ (PhyloHMM.hpp:27)   */

  int32_t has_d;             /* 1 = igh (7 regions), 0 = igk/igl (5 regions)  */
  int32_t n_v, n_d, n_j;     /* germline states per region (= genes)          */
  int32_t vd_sites, vd_states; /* V-D (or V-J) junction: L x S                */
  int32_t dj_sites, dj_states; /* D-J junction; 0 x 0 when !has_d             */

  /* germline / padding regions: CSR lists of xMSA columns per gene, gene order
   * = std::map order of *_ggene_ranges_ (src/PhyloHMM.cpp:158-193)           */
  const int32_t* vpad_off;  const int32_t* vpad_inds;   /* off[n_v+1]         */
  const int32_t* vgerm_off; const int32_t* vgerm_inds;  /* off[n_v+1]         */
  const int32_t* dgerm_off; const int32_t* dgerm_inds;  /* off[n_d+1]         */
  const int32_t* jgerm_off; const int32_t* jgerm_inds;  /* off[n_j+1]         */
  const int32_t* jpad_off;  const int32_t* jpad_inds;   /* off[n_j+1]         */

  /* junction xMSA index matrices, column-major [sites x states], -1 = no
   * emission (src/PhyloHMM.cpp:202-215, 489-513)                             */
  const int32_t* vd_inds;
  const int32_t* dj_inds;

  /* V-germline start terms (src/HMM.cpp:307-314)                             */
  const double* v_gene_prob;   /* [n_v] germ_ptr->gene_prob()                 */
  const double* v_trans_prod;  /* [n_v] prod of in-region germline transitions*/
  const double* vpad_trans;    /* [n_v] vpadding_transition_                  */
  const double* jpad_trans;    /* [n_j] jpadding_transition_                  */

  /* transition matrices, column-major (src/HMM.hpp transition members)       */
  const double* t_v_vd;   /* [n_v x vd_states]       vgerm_vd_junction_       */
  const double* t_vd;     /* [vd_states x vd_states] vd_junction_             */
  const double* t_vd_d;   /* [vd_states x n_d] (n_j when !has_d) vd_junction_dgerm_ */
  const double* t_d_dj;   /* [n_d x dj_states]       dgerm_dj_junction_       */
  const double* t_dj;     /* [dj_states x dj_states] dj_junction_             */
  const double* t_dj_j;   /* [dj_states x n_j]       dj_junction_jgerm_       */
} lh_family_desc_t;

/* B tree samples in structure-of-arrays form; replaces pll_utree_t +
 * pll_operation_t (libpll pll.h:249-259, 312-334) and pt::pll::Model
 * (lib/libptpll/src/model.hpp:10-16).
 * Node ids: tips 0..T-1 are rows of tip_codes; inner node T+i is produced by
 * operation i. Operations are in any valid post-order (children before
 * parents); operation T-3 is the virtual root (pll_util.cpp:368-371), whose
 * remaining neighbour across the evaluated edge is root_child.               */
typedef struct lh_tree_batch {
  int32_t n_trees;            /* B                                            */
  int32_t n_rates;            /* R, 1..LH_MAX_RATES                           */
  const int32_t* children;    /* [B][T-2][2]                                  */
  const double*  blen;        /* [B][T-2][2] branch above each child          */
  const int32_t* root_child;  /* [B]                                          */
  const double*  root_blen;   /* [B]                                          */
  const double*  er;          /* [B][6] AC AG AT CG CT GT (models.c:225-237)  */
  const double*  pi;          /* [B][4] A C G T                               */
  const double*  rates;       /* [B][R] discrete-Gamma category rates         */
} lh_tree_batch_t;

/* B tree samples as the RevBayes trace holds them: Newick text plus the substitution parameters of the
 * row (src/PhyloHMM.cpp:414-426). The strings go to the device as they are; parsing
 * (regex_replace("[&index=N]") + pll_utree_parse_newick_string + set_missing_branch_length + the
 * post-order traversal: src/PhyloHMM.cpp:419-422, parse_utree.y:179-246, pll_util.cpp:46-62,
 * utree.c:284-442) is a kernel there. Tip labels are matched against the family's xMSA row labels
 * (lh_family_set_labels), as Partition::SetTipStates matches them (pll_partition.cpp:153-194).     */
typedef struct lh_newick_batch {
  int32_t n_trees;            /* B                                                       */
  int32_t n_rates;            /* R                                                       */
  const char* const* newick;  /* [B] NUL-terminated Newick strings                       */
  const int64_t* length;      /* [B] strlen of each string, or NULL                      */
  const double*  er;          /* [B][6]                                                  */
  const double*  pi;          /* [B][4]                                                  */
  const double*  rates;       /* [B][R] category rates, or NULL: computed from alpha     */
  const double*  alpha;       /* [B] Gamma shape (pll_compute_gamma_cats, MEAN); read when rates is NULL */
} lh_newick_batch_t;

/* Optional per-tree read-outs (any pointer may be NULL). What the reference
 * exposes through accessors after LogLikelihood(): xmsa_emission()
 * (PhyloHMM.hpp:31), the forward tables and scaler counts (HMM.hpp forward
 * members) needed by SampleNaiveSequence (src/HMM.cpp:358-431).
 * Junction tables are column-major [sites x states] per tree.                */
typedef struct lh_outputs {
  double*  xmsa_emission;   /* [B][S_x]                                       */
  double*  vgerm_forward;   /* [B][n_v]                                       */
  double*  vd_forward;      /* [B][vd_sites*vd_states]                        */
  double*  dgerm_forward;   /* [B][n_d]                                       */
  double*  dj_forward;      /* [B][dj_sites*dj_states]                        */
  double*  jgerm_forward;   /* [B][n_j]                                       */
  int32_t* germ_scalers;    /* [B][3] vgerm, dgerm, jgerm scaler counts       */
  int32_t* vd_scalers;      /* [B][vd_sites]                                  */
  int32_t* dj_scalers;      /* [B][dj_sites]                                  */
  /* per rate category k and xMSA column: log( L_k(column) / pi[naive base] ), the unweighted category
   * likelihoods whose mean is xmsa_emission. This is what the reference's ancestral-sequence step
   * recomputes per kept tree, site and rate with phylomd (scripts/run_bootstrap_asr_ess.R:84-87) to
   * sample the rate category of a site; -inf where the category's likelihood is not positive.       */
  double*  site_rate_loglik; /* [B][R][S_x]                                   */
} lh_outputs_t;

typedef struct lh_timings {
  float ms_h2d;              /* host->device copies of the batch (CUDA events)           */
  float ms_plan;             /* host traversal planning (wall clock)                     */
  float ms_pmatrix;          /* sum over k_pmatrix launches                              */
  float ms_prune;            /* sum over k_prune launches                                */
  float ms_edge;             /* sum over k_edge launches                                 */
  float ms_hmm;              /* sum over k_hmm launches                                  */
  float ms_d2h, ms_total;    /* device->host; first H2D event to last D2H event          */
  float ms_kernels;          /* k_eigen start to k_hmm end                               */
  int32_t n_trees;
  int32_t kernel_launches;   /* launches of this library's kernels                       */
  int32_t prune_launches;    /* of which k_prune                                         */
  int64_t h2d_bytes, d2h_bytes;
  int32_t prune_slots;       /* CLV stack depth the batch needed                         */
  int32_t prune_smem_slots;  /* of which held in shared memory                           */
  int32_t prune_smem_lo;     /* first stack level of the shared-memory window            */
  int32_t prune_tile_sites;  /* xMSA columns per CTA                                     */
  int32_t prune_sites_per_thread;
  int32_t prune_grid, prune_ctas_per_sm;
  int32_t prune_columns;     /* columns k_prune walked: S_x, or the distinct site patterns */
  int32_t pattern_trees;     /* trees of the batch pruned on site patterns               */
  float ms_call;             /* wall clock of the last lh_score_batch call               */
  float ms_setup;            /* of which validation + buffer set-up before the pipeline  */
  float ms_plan_dev;         /* sum over k_plan launches (device traversal planning)     */
  int32_t device_plan;       /* 1 when the batch was planned by k_plan                   */
  float ms_newick;           /* sum over k_newick launches (device Newick parsing)       */
  int32_t newick_host_trees; /* trees of an lh_score_newick call the host parser took over */
  int32_t text_direct;       /* chunks of an lh_score_newick call whose text was copied in place from the caller's
                              * pinned memory (one 2-D copy) instead of being staged through the pinned ring   */
} lh_timings_t;

const char* lh_last_error(void);
const char* lh_version(void);
int lh_device_count(int* n);

int  lh_family_create(lh_family_t** out, const lh_family_desc_t* desc, int device);
void lh_family_destroy(lh_family_t* fam);
/* Bind another clonal family to an existing handle (what constructing the next PhyloHMM object,
 * src/PhyloHMM.cpp:17-40, is to the reference): the family constants go to the handle's device
 * arena in one copy; streams, planner threads, pinned staging and batch buffers are reused.
 * After a failed rebind the handle holds no family until a rebind succeeds.                     */
int  lh_family_rebind(lh_family_t* fam, const lh_family_desc_t* desc);

/* Score B samples: host buffers in, host log-likelihoods out (synchronous).
 * Includes the host->device and device->host copies.                         */
int lh_score_batch(lh_family_t* fam, const lh_tree_batch_t* trees,
                   double* loglik /* [B] */, lh_outputs_t* opt /* or NULL */);

/* The same for samples given as Newick text (the batched RevBayes hook: what the loop body
 * src/PhyloHMM.cpp:419-430 does per row, for B rows). Needs lh_family_set_labels first. A tree the
 * device parser does not accept is parsed by the host parser (same grammar as the reference's); a
 * malformed tree fails the call with that parser's message.                                      */
int lh_family_set_labels(lh_family_t* fam, const char* const* labels /* [n_tips], xMSA row order */);
int lh_score_newick(lh_family_t* fam, const lh_newick_batch_t* trees,
                    double* loglik /* [B] */, lh_outputs_t* opt /* or NULL */);

/* Split form for callers that keep a batch resident (bench `value`, pipelining):
 * upload once, run any number of times, fetch.                               */
int lh_batch_upload(lh_family_t* fam, const lh_tree_batch_t* trees);
int lh_batch_run(lh_family_t* fam);                  /* async on the family stream */
int lh_batch_fetch(lh_family_t* fam, double* loglik, lh_outputs_t* opt); /* syncs */
int lh_sync(lh_family_t* fam);
int lh_loglik_dev(lh_family_t* fam, const double** loglik_dev, int32_t* n); /* device ptr of the last run */

/* Ancestral-state sampling: the second half of the reference's R step
 * (scripts/run_bootstrap_asr_ess.R:81-89 - per kept tree and alignment site, the rate category drawn in proportion
 * to phylomd::phylo.likelihood on the rate-scaled tree, then phylomd::asr.sim on that tree:
 * lib/phylomd/src/simulation_functions.cpp:57-117). The alignment is the family's MSA - not the xMSA - with the
 * "naive" row replaced by the sample's NaiveSequence (R:66-67): msa_codes holds the rows in the family's tip order
 * (codes A,C,G,T = 0..3, 4 = N / gap), row naive_row of it is ignored and read from naive_codes per tree
 * (naive_row = -1: no row is replaced). Trees and substitution parameters as for lh_score_batch.
 * Outputs: node_states [B][T-2][n_sites], the sampled base (0..3; 4 where the site has no positive likelihood) of
 * inner node T + op for every operation op of the tree (the last one is the trifurcation next to root_child, which is
 * the script's root when the tree is rooted at the naive outgroup); site_rate [B][n_sites], the sampled category
 * (n_rates where none has a positive likelihood); site_rate_loglik [B][R][n_sites] or NULL,
 * log( L_k(site) / pi[naive base] ) as R:85 forms it (pi = 1 for an N).
 * Draws come from Philox4x32-10 keyed by `seed` with counter (tree, site, draw): the result depends on nothing but
 * the arguments. The reference draws with R's generator on a worker cluster, which nothing outside R reproduces;
 * parity with it is distributional, parity with oracle/asr_oracle.py exact (tests/test_gpu_asr.py). Synchronous;
 * keeps one device pool per handle between calls (option "asr_scratch_mb": CLV scratch per chunk of trees, default 4096).   */
typedef struct lh_asr_batch {
  const lh_tree_batch_t* trees;
  int32_t n_sites;
  int32_t naive_row;
  const uint8_t* msa_codes;    /* [T][n_sites]                                           */
  const uint8_t* naive_codes;  /* [B][n_sites]                                           */
  uint64_t seed;
} lh_asr_batch_t;
int lh_asr_sample(lh_family_t* fam, const lh_asr_batch_t* batch, uint8_t* node_states, uint8_t* site_rate,
                  double* site_rate_loglik);

/* The random draws of HMM::SampleNaiveSequence (src/HMM.cpp:323-341, 1222-1353) for every tree of the batch the
 * handle scored last, on the device: the J gene, the D-J junction states from the last site to the first, the D gene,
 * the V-D junction states, the V gene - each a std::discrete_distribution draw over (transition column of the state
 * drawn before) x (forward row), reproduced operation by operation as libstdc++ computes it. The forward tables must
 * still be on the device: set option "keep_tables" = 1 before the scoring call (or request them as read-outs).
 * words [B][n_words]: the raw 32-bit outputs of the run's std::mt19937 in the order the draws consume them - two per
 * draw over two or more weights, so n_words = 2 * (1 + dj_sites + 1 + vd_sites + 1) for a heavy chain with more than
 * one candidate everywhere; the call checks the count. idx [B][n_idx]: J gene | D-J states (site 0 first) | D gene |
 * V-D states | V gene (light chains: J gene | V-J states | V gene). flags [B]: 1 where the tree must be sampled by
 * the host code instead (a non-finite forward entry, weights that do not sum to a positive finite number).        */
int lh_sample_paths(lh_family_t* fam, int32_t n_trees, const uint32_t* words, int32_t n_words, int32_t* idx,
                    int32_t n_idx, uint8_t* flags);

/* Page-lock a range of the caller's own memory for every device of the process (cudaHostRegister). Newick text
 * that lh_score_newick finds in pinned memory, with its lengths given and one string every constant number of bytes
 * (8-byte aligned), is copied to the device where it lies - one 2-D copy per chunk, no staging pass of the host over
 * the bytes (lh_timings_t.text_direct counts those chunks; option "text_direct" = 0 turns it off). Memory from
 * cudaHostAlloc / torch pin_memory qualifies as it is.                                                         */
int lh_pin_host(void* ptr, size_t bytes);
int lh_unpin_host(void* ptr);
int lh_get_timings(lh_family_t* fam, lh_timings_t* out); /* of the last run/score */
/* Test hook: the traversal programs of the current batch as the device holds them
 * (steps [B][T-1][4] int32, pblen [B][2T-3]); compared with the host planner in the tests. */
int lh_debug_fetch_plan(lh_family_t* fam, int32_t* steps, double* pblen);
/* Test hook: runs only the device Newick parser (k_newick) over the batch's strings and returns what it
 * wrote - children [B][T-2][2], blen [B][T-2][2], root_child [B], root_blen [B], whether the naive tip is on
 * the root edge or under the virtual root, and whether the tree was left to the host parser - plus the
 * summed kernel time; compared with lhh_newick_to_ops in the tests.                                      */
int lh_debug_parse_newick(lh_family_t* fam, const lh_newick_batch_t* trees, int32_t* children, double* blen,
                          int32_t* root_child, double* root_blen, uint8_t* on_patterns /* or NULL */,
                          uint8_t* left_to_host, float* ms_kernel /* or NULL */);
/* The device Newick parser (k_newick_cta + k_newick) for callers that cannot allocate - code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc. This entry point has NO counterpart in the
 * reference. It calls no cudaMalloc / cudaFree or any other
 * allocating routine: the parser's per-tree scratch lives in a static device array of
 * LH_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit, and all inputs and outputs are
 * the caller's DEVICE buffers:
 *   text_dev   the strings, each NUL-terminated, starting on a 16-byte boundary and padded with NULs to one
 *   offs_dev   [n_trees] start of each string in text_dev;  lens_dev [n_trees] its strlen
 *   children_dev [n][T-2][2], blen_dev [n][T-2][2], root_child_dev [n], root_blen_dev [n]  as lh_tree_batch_t
 *   on_patterns_dev [n] 1: naive tip on the root edge or under the virtual root
 *   status_dev [n] 0 parsed; non-zero: not decided here (malformed, or a spelling left to the host parser) -
 *              the tree's rows then hold a harmless stand-in
 * The family needs lh_family_set_labels (set-up; that call does allocate). The two kernels are queued on `stream`
 * (a cudaStream_t; NULL = the legacy default stream) and the call returns without waiting. Limit: n_trees *
 * ((2T-2)*14 rounded up to 16) bytes <= LH_NOALLOC_SCRATCH_BYTES, T < 32768 - lh_newick_parse_noalloc_max_trees(T)
 * trees per call (466 for 1001 tips); more is refused with an error. Every call shares the one array per GPU: it
 * keeps the last call's contents (the kernel writes every item before reading it, so nothing needs clearing), and
 * calls that overlap on one GPU - two streams, two threads, two handles - race on it.                         */
#define LH_NOALLOC_SCRATCH_BYTES (16u << 20)
int lh_newick_parse_noalloc(lh_family_t* fam, int32_t n_trees, const char* text_dev, const uint64_t* offs_dev,
                            const uint32_t* lens_dev, int32_t* children_dev, double* blen_dev, int32_t* root_child_dev,
                            double* root_blen_dev, uint8_t* on_patterns_dev, uint8_t* status_dev, void* stream);
int32_t lh_newick_parse_noalloc_max_trees(int32_t n_tips);
/* Diagnostic: sustained FP64 FMA rate of a register-resident DFMA loop (TFLOP/s); the
 * denominator for the FP64-pipe view of the pruning kernel (no reference counterpart). */
int lh_measure_fp64_peak(int device, double* tflops);
int lh_set_option(lh_family_t* fam, const char* key, int64_t value);
/* keys: "sites_per_thread" (0 = auto, 1..4: xMSA columns per thread; a CTA is one warp and
 * covers 32x that many columns), "smem_slots" (-1 = auto), "smem_lo" (-1 = auto: first CLV stack level
 * of the shared-memory window), "timing" (0/1), "host_threads", "hmm_overlap" (1 = default: the forward
 * pass of a chunk runs on its own stream beside the next chunk's pruning),
 * "chunk_growth" (1 = default: later chunks of the pipelined call grow towards the resident batch's size
 * while the host stages faster than the GPU scores),
 * "open_unit" (1 = default: the pipelined call opens with a chunk of one work item per resident CTA),
 * "device_plan" (1: the traversal programs are built on the device by k_plan from the operation
 * lists as the caller holds them; 0: by the host planner threads; -1 = default: by the device when
 * the handle's host threads would not keep ahead of the kernels - the programs are the same),
 * "newick_kernel" (lh_score_newick: 0 / 2 = default, the one-CTA-per-tree parser, with the one-thread-per-tree
 * parser behind it for trees nested deeper than its tables; 3 = the one-warp-per-tree parser (ballot lexing) in its
 * place; 1 = the one-thread-per-tree parser alone - the three write the same bits),
 * "parse_reserve" (default 0: pruning CTAs per SM given up during lh_score_newick so that the parser of the
 * next chunk runs beside the pruning of this one),
 * "chunk_bytes" (branch-record bytes per k_pmatrix/k_prune launch pair; the chunk is never
 * smaller than 16 waves of CTAs), "chunk_trees" (0 = auto; trees per chunk, overrides), "pipe_waves" (default 4:
 * lh_score_batch launches a chunk's kernels every that many waves of CTAs), "stream" (a
 * caller-owned cudaStream_t cast to int64 - 0 is the legacy default stream, i.e. torch's default
 * stream - on which the kernels and lh_loglik_dev's buffer are then ordered; -1 = back to a private stream), "site_patterns" (1 =
 * default: trees that have the naive tip on the root edge or under the virtual root are
 * pruned on the distinct site patterns of xMSA rows 1.., and k_edge applies the naive row's
 * factor per xMSA column; other trees, and every tree when 0, are pruned on all S_x
 * columns. The results are the same doubles either way)                                */

/* CUDA-event stopwatch on the family's stream (8 slots), so callers can time a region of
 * lh_batch_run calls on the stream the kernels are launched on.                          */
int lh_timer_mark(lh_family_t* fam, int32_t slot);
int lh_timer_elapsed_ms(lh_family_t* fam, int32_t from_slot, int32_t to_slot, float* ms);

/* ---- group: one clonal family on several GPUs of one box, inside one process ----
 * The reference has no counterpart (one CPU process per cluster). The trees of a call are cut into contiguous,
 * equal shards, one per device; the family constants are replicated; each member is an lh_family_t driven by
 * its own host thread; results are written into the caller's single host arrays. The doubles do not depend on
 * the number of devices. A group of one device is an lh_family_t with another name.                       */
typedef struct lh_group lh_group_t;
int  lh_group_create(lh_group_t** out, const lh_family_desc_t* desc, const int* devices, int n_devices);
void lh_group_destroy(lh_group_t* grp);
int  lh_group_size(const lh_group_t* grp);
lh_family_t* lh_group_member(lh_group_t* grp, int member);            /* owned by the group */
int  lh_group_set_labels(lh_group_t* grp, const char* const* labels);
int  lh_group_set_option(lh_group_t* grp, const char* key, int64_t value);   /* lh_set_option on every member */
int  lh_group_score_batch(lh_group_t* grp, const lh_tree_batch_t* trees, double* loglik, lh_outputs_t* opt);
int  lh_group_score_newick(lh_group_t* grp, const lh_newick_batch_t* trees, double* loglik, lh_outputs_t* opt);
int  lh_group_get_timings(lh_group_t* grp, int member /* -1: summed counts, slowest member's times */, lh_timings_t* out);
const char* lh_group_last_error(void);
/* "0-3", "0,2,5", "1" -> device ordinals: the CLI's --devices flag and the LINEARHAM_DEVICES variable */
int  lh_parse_devices(const char* spec, int* out, int cap, int* n_out);

/* ---- repertoire: many independent clonal families on one device (BASELINE config 5) --------
 * The reference runs one linearham process per cluster (SConstruct:344-352, 428-447;
 * src/linearham.cpp:377-386). A repertoire keeps n_slots device handles; one host thread per slot
 * pulls families off a queue (largest first), rebinds its handle and calls lh_score_batch, so the
 * host planning / copies of one family overlap the kernels of another. Results do not depend on
 * the slot, the order or n_slots.                                                             */
typedef struct lh_repertoire lh_repertoire_t;
typedef struct lh_repertoire_item {
  const lh_family_desc_t* desc;   /* the family                                          */
  const lh_tree_batch_t*  trees;  /* its tree samples                                    */
  double*  loglik;                /* out [trees->n_trees]                                */
  int32_t  status;                /* out: 0 scored, 1 failed, -1 not reached             */
} lh_repertoire_item_t;
typedef struct lh_repertoire_stats {
  float   ms_wall;                /* wall clock of the call                              */
  float   ms_bind, ms_score;      /* summed over slots: create/rebind, lh_score_batch    */
  int32_t n_families, n_failed, n_slots, host_threads_per_slot;
  int64_t n_trees;                /* scored                                              */
  int64_t kernel_launches, h2d_bytes, d2h_bytes;
} lh_repertoire_stats_t;
int  lh_repertoire_create(lh_repertoire_t** out, int device, int n_slots, int host_threads /*0 = cores / n_slots*/);
/* the same over several devices of the box: n_slots handles on each, ONE queue (largest family first) that
 * every slot of every device pulls from                                                          */
int  lh_repertoire_create_multi(lh_repertoire_t** out, const int* devices, int n_devices, int n_slots,
                                int host_threads /*0 = cores / (n_devices * n_slots)*/);
int  lh_repertoire_score(lh_repertoire_t* rep, lh_repertoire_item_t* items, int32_t n_items,
                         lh_repertoire_stats_t* stats /* or NULL */);
int  lh_repertoire_set_option(lh_repertoire_t* rep, const char* key, int64_t value); /* lh_set_option on every slot */
void lh_repertoire_destroy(lh_repertoire_t* rep);
const char* lh_repertoire_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LINEARHAM_B200_H_ */
