// Device-side data layout and kernel launchers (sm_100a). See DESIGN.md.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>

namespace lhb {

constexpr int kEigStride = 44;  // per tree: evals[4] evecs[16] inv_evecs[16] log(pi)[4] pi[4]

// Sparse column form of one HMM transition matrix (family constant).
struct CscDev {
  const int32_t* colptr;  // [cols+1]
  const double2* ent;     // [nnz] x = value, low word of y = row (one 16-byte load per non-zero)
  int32_t rows, cols;
};

struct FamilyDev {
  int32_t T, S;                 // tips, xMSA columns
  const uint8_t* tip_codes;     // [T][S]
  int32_t has_d, n_v, n_d, n_j;
  int32_t vd_sites, vd_states, dj_sites, dj_states;
  const int32_t *vpad_off, *vpad_inds, *vgerm_off, *vgerm_inds, *dgerm_off, *dgerm_inds;
  const int32_t *jgerm_off, *jgerm_inds, *jpad_off, *jpad_inds;
  const int32_t* vd_inds_rm;    // [vd_sites][vd_states] row-major copy
  const int32_t* dj_inds_rm;
  // the states that can emit at a junction site (xMSA index >= 0), site by site:
  // pairs (state, xMSA index) [valid_off[i], valid_off[i+1]) for site i
  const int32_t *vd_valid_off, *dj_valid_off;
  const int2 *vd_valid, *dj_valid;
  int32_t skip_silent;          // 1: every transition entry is small enough for the skip below to be exact
  const double *v_gene_prob, *v_trans_prod, *vpad_trans, *jpad_trans;
  CscDev t_v_vd, t_vd, t_vd_d, t_d_dj, t_dj, t_dj_j;
  int32_t g_max, s_max;         // max genes per region, max states per region
};

// Branch records of one (tree, category) pass: the 2T-3 branches in the order the program consumes
// them, then one identity record (the root CLV's side of the edge evaluation).
#ifdef __CUDACC__
__host__ __device__
#endif
inline int records_per_pass(int n_tips) { return 2 * n_tips - 2; }

struct BatchDev {
  int32_t B, R;
  const int4* plan;      // [B][T-1]
  const double* pblen;   // [B][2T-3]
  const double* er;      // [B][6]
  const double* pi;      // [B][4]
  const double* rates;   // [B][R]
  double* eig;           // [B][kEigStride]
  double* prec;          // [chunk][R][records_per_pass(T)][20] branch records
  double* xe;            // [B][S] xmsa_emission
  double* loglik;        // [B]
};

struct PruneCfg {
  int32_t sites_per_thread;  // NS; a CTA (one warp) covers 32*NS xMSA columns
  int32_t n_tiles;
  int32_t smem_slots;    // stack slots kept in shared memory ...
  int32_t smem_lo;       // ... starting with this one (the host picks the busiest window)
  int32_t total_slots;   // stack slots the batch needs
  int32_t grid;          // persistent CTAs
  double* scratch;       // [grid][total_slots-smem_slots][4 NS][32] doubles
  int32_t* scratch_cnt;  // [grid][total_slots-smem_slots][NS][32]
  double log_threshold;  // log(2^-256), host libm (core_likelihood_avx.c:1249)
  // tip codes re-tiled for this NS: word [tip][tile][lane], byte u = 32 * code of column
  // tile*32*NS + u*32 + lane (0 beyond S); 32*code is the byte offset of the Q column
  const uint32_t* codes_t;
  // pruning columns. k_prune walks `n_cols` columns: either all S xMSA columns, or - for trees
  // that have the naive tip (xMSA row 0) on the root edge or under the virtual root - the U
  // distinct site patterns of rows 1..T-1, with the naive tip's factor deferred to k_edge.
  int32_t n_cols;        // S or U
  int32_t cols_pad;      // n_tiles * 32 * NS
  int32_t defer_naive;   // 1: pattern columns, naive factor applied by k_edge
  const int32_t* tree_list;  // trees of this launch (batch indices inside the chunk)
  // Work items beyond a CTA's first are handed out through this counter (zeroed before the launch):
  // a CTA that starts late - the forward pass of the previous chunk may still hold part of an SM -
  // simply takes fewer items.
  int32_t* work_counter;
  // Newick input: which trees of the chunk run on patterns is decided on the device (k_newick, k_split).
  // tree_list then is the chunk's whole list (pattern trees first) and *dyn_n_pat how many of its
  // dyn_total trees are pattern trees; the launch with defer_naive takes those, the other one the rest.
  const int32_t* dyn_n_pat;
  int32_t dyn_total;
  // per (chunk tree, category, column) inputs of the edge term, written by k_prune:
  double* ebuf;          // tree at e_stride doubles: [R][8][cols_pad] message across the root edge [4], root CLV [4]
  int32_t* ecnt;         // tree at c_stride ints:    [R][cols_pad] scaler count of the two
  int32_t* enaive;       // [chunk] record index of the naive tip's branch, -1 when not deferred
  int64_t e_stride, c_stride;
};

// k_edge reads what the k_prune launches of a chunk left behind
struct EdgeCfg {
  const double* ebuf;
  const int32_t* ecnt;
  const int32_t* enaive;
  int64_t e_stride, c_stride;
  int32_t cols_pad_all, cols_pad_pat;  // column pitch of trees pruned on all columns / on patterns
  const int32_t* col_of_site;          // [S] pattern index of each xMSA column
  double log_threshold;
  double* rate_ll;                     // optional [B][R][S]: per-category log( L_k / pi[naive] )
};

struct HmmOut {           // optional device buffers (may be null)
  double *vgerm_forward, *vd_forward, *dgerm_forward, *dj_forward, *jgerm_forward;
  int32_t *germ_scalers, *vd_scalers, *dj_scalers;
  double log_factor;      // log(2^256), host libm (src/HMM.cpp:353)
};

// k_plan (plan_dev.cu): the traversal planner on the device; arrays are indexed by batch tree
struct PlanDev {
  int32_t T;
  int32_t tree0;               // first tree of the launch
  const int32_t* children;     // [B][T-2][2] as in lh_tree_batch_t
  const double* blen;          // [B][T-2][2]
  const int32_t* root_child;   // [B]
  const double* root_blen;     // [B]
  int4* plan;                  // [B][T-1] out
  double* pblen;               // [B][2T-3] out
  int32_t* flags;              // [0] min tree with a negative branch length, [1] min malformed tree
                               // (both INT_MAX when none), [2] deepest CLV stack met
};
// k_newick (newick_dev.cu): Newick text -> the operation arrays k_plan consumes, one thread per tree.
// Tip labels are found through an open-addressing table on the FNV-1a hash of the label:
//   slot = { first 8 label bytes (little-endian, zero-padded) lo, hi; hash; length | row << 16 },
//   length | row << 16 == 0xffffffff marks an empty slot; labels longer than 8 bytes are compared
//   against chars + long_off[row] as well.
struct LabelTableDev {
  const uint4* slots;
  uint32_t mask;               // slot count - 1 (power of two)
  const uint32_t* long_off;    // [T] offset of each label's bytes in chars
  const char* chars;
};
struct NewickDev {
  int32_t T;
  int32_t tree0, n;            // batch index of the chunk's first tree, trees in the chunk
  const char* text;            // the chunk's strings, each NUL-terminated and starting on a 16-byte boundary
  const uint64_t* offs;        // [n] start of each string in text
  const uint32_t* lens;        // [n] strlen of each string
  int32_t kernel;              // 0 / 2: k_newick_cta (one CTA per tree), then k_newick for the trees too deep for
                               //        its tables; 1: k_newick alone (one thread per tree)
  unsigned char* items;        // k_newick_cta scratch: newick_items_bytes(T, n)
  int32_t only_deferred;       // (set by launch_newick) k_newick takes only the trees k_newick_cta marked status 2
  LabelTableDev labels;
  int32_t* children;           // [B][T-2][2]   (batch arrays, as in PlanDev)
  double* blen;                // [B][T-2][2]
  int32_t* root_child;         // [B]
  double* root_blen;           // [B]
  uint8_t* on_pat;             // [B] 1: the naive tip (row 0) is on the root edge or under the virtual root
  uint8_t* status;             // [B] 0 parsed; 1 left to the host parser (malformed, or a spelling / number
                               //     this parser does not decide) - the tree then gets a harmless stand-in
};
void launch_newick(const NewickDev& p, cudaStream_t s);
size_t newick_items_bytes(int n_tips, int n_trees);
// the no-allocation variant: k_newick_cta's scratch in a static device array of this size (newick_dev.cu)
constexpr size_t kNoAllocScratchBytes = (size_t)16 << 20;
int newick_noalloc_max_trees(int n_tips);   // trees per call the static scratch holds (0: family too large)
int launch_newick_noalloc(NewickDev p, int device, cudaStream_t s);   // 0 or a cudaError_t
// stable partition of the chunk's trees [t0, t0+n) by on_pat into list[t0..) (pattern trees first);
// n_pat_out[0] = how many
void launch_split(const uint8_t* on_pat, int t0, int n, int32_t* list, int32_t* n_pat_out, cudaStream_t s);

// Raises a kernel's dynamic shared-memory limit to at least `bytes`; the limit only ever grows, so
// handles that launch the same kernel with different sizes from different host threads do not race.
void ensure_dynamic_smem(const void* kernel, size_t bytes);
size_t plan_smem_bytes(int n_tips);
bool plan_on_device_ok(int n_tips);
void launch_plan(const PlanDev& p, int ntrees, cudaStream_t s);

void launch_eigen(const BatchDev& b, int tree0, int ntrees, cudaStream_t s);
void launch_pmatrix(const FamilyDev& f, const BatchDev& b, int tree0, int ntrees, cudaStream_t s);
// prunes the `nlist` trees c.tree_list names; tree0 is the first tree of the chunk (its branch
// records and edge buffers are chunk-relative)
void launch_prune(const FamilyDev& f, const BatchDev& b, const PruneCfg& c, int tree0, int nlist,
                  cudaStream_t s);
void launch_edge(const FamilyDev& f, const BatchDev& b, const EdgeCfg& c, int tree0, int ntrees,
                 cudaStream_t s);
void launch_hmm(const FamilyDev& f, const BatchDev& b, const HmmOut& o, int tree0, int ntrees,
                cudaStream_t s);
size_t prune_smem_bytes(int sites_per_thread, int smem_slots);
int prune_max_ctas_per_sm(int sites_per_thread, int smem_slots);
size_t hmm_smem_bytes(const FamilyDev& f, int warps);
// launches the DFMA diagnostic kernel; returns the flop count of the launch
double fp64_peak_flops_per_launch(double* scratch, int blocks, int iters, cudaStream_t s);

// k_sample: the index draws of SampleNaiveSequence for every tree of a scored batch (forward tables on the device)
void launch_sample(const FamilyDev& f, const double* fv, const double* fvd, const double* fd, const double* fdj,
                   const double* fj, const uint32_t* words, int n_words, int32_t* idx, int n_idx, uint8_t* flag,
                   int ntrees, cudaStream_t s);

// asr.cu: ancestral-state sampling (lh_asr_sample); synchronous; *cache keeps its device pool between calls
// (nullptr at first, released with asr_cache_free). 0, or 1 with *err set.
void asr_cache_free(void* cache);
int asr_sample(void** cache, int n_tips, int n_trees, int n_rates, const int32_t* children, const double* blen,
               const int32_t* root_child, const double* root_blen, const double* er, const double* pi,
               const double* rates, int n_sites, const uint8_t* msa, int naive_row, const uint8_t* naive,
               uint64_t seed, size_t scratch_bytes, uint8_t* node_states, uint8_t* site_rate, double* site_rate_loglik,
               std::string* err);

}  // namespace lhb
