// Host-side mirror of the reference's HMM / PhyloHMM classes for the phylo-HMM likelihood path
// (src/HMM.hpp:31-413, src/PhyloHMM.hpp:24-127): same constructor arguments, same method names
// and meaning, same error behaviour (std::runtime_error / std::invalid_argument), so that code
// and tests written against the reference read the same. What differs is underneath:
//   * everything that runs per tree sample (tip expansion, pruning, emission fill, forward pass)
//     is one call into the CUDA library (include/linearham_b200.h); there is no CPU path
//   * Eigen is not used: matrices are plain column-major buffers (Matrix<T>), which is also the
//     layout the device ABI takes
//   * RunPipeline scores the RevBayes samples in batches (new third argument), then samples a
//     naive sequence per row on the host exactly as the reference does
// Construction (YAML -> state space -> transition matrices -> xMSA) happens once per family and
// stays on the host; it must be bit-exact.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <utility>
#include <vector>

#include "germline.hpp"
#include "yaml_lite.hpp"

struct lh_family;
struct lh_group;       // include/linearham_b200.h
struct lh_family_desc;  // include/linearham_b200.h

namespace linearham {

// Column-major dense matrix (the layout of Eigen::Matrix<T, Dynamic, Dynamic>).
template <class T>
struct Matrix {
  int r = 0, c = 0;
  std::vector<T> v;
  void assign(int rows, int cols, T fill) {
    r = rows;
    c = cols;
    v.assign((std::size_t)rows * cols, fill);
  }
  int rows() const { return r; }
  int cols() const { return c; }
  std::size_t size() const { return v.size(); }
  T& operator()(int i, int j) { return v[(std::size_t)j * r + i]; }
  const T& operator()(int i, int j) const { return v[(std::size_t)j * r + i]; }
  const T* data() const { return v.data(); }
  T* data() { return v.data(); }
  bool operator==(const Matrix& o) const { return r == o.r && c == o.c && v == o.v; }
};
typedef Matrix<double> MatrixXd;
typedef Matrix<int> MatrixXi;
typedef std::map<std::string, std::pair<int, int>> GeneRanges;

// One "germline" or "padding" region: one HMM state per candidate gene (src/HMM.hpp:49-113).
struct GermlineRegion {
  std::vector<std::string> state_strs;
  std::vector<int> left_del, right_del;
  GeneRanges ggene_ranges;
  std::vector<int> naive_bases, germ_inds, site_inds;
  std::vector<int> xmsa_inds;              // PhyloHMM: xMSA column of every (state, site)
  std::vector<double> forward;             // [states] after a forward pass
  int scaler_count = 0;
};

// One "junction" region (src/HMM.hpp:62-98).
struct JunctionRegion {
  std::vector<std::string> state_strs;
  std::vector<int> del;
  std::vector<char> ggene_types;           // 'V' / 'D' / 'J'
  GeneRanges ggene_ranges;
  std::vector<int> naive_bases, germ_inds, site_inds;
  MatrixXi xmsa_inds;                      // [sites x states], -1 = state cannot emit at the site
  MatrixXd forward;                        // [sites x states]
  std::vector<int> scaler_counts;          // [sites]
};

class HMM {
 public:
  HMM(const std::string& yaml_path, int cluster_ind, const std::string& hmm_param_dir, int seed);
  virtual ~HMM() {}

  virtual double LogLikelihood() = 0;
  std::string SampleNaiveSequence();
  // the same for a path whose states were drawn on the device (lh_sample_paths): no forward table is read, no random
  // number drawn; idx as that call lays it out
  std::string ApplySampledPath(const int32_t* idx);

  // ---- accessors (names as in src/HMM.hpp:197-409) ------------------------------------------
  const std::string& locus() const { return locus_; }
  const std::map<std::string, std::pair<int, int>>& flexbounds() const { return flexbounds_; }
  const std::map<std::string, int>& relpos() const { return relpos_; }
  const std::map<std::string, lhb::GermlineGene>& ggenes() const { return ggenes_; }
  const std::string& alphabet() const { return alphabet_; }
  const MatrixXi& msa() const { return msa_; }
  const GermlineRegion& vpadding() const { return vpadding_; }
  const GermlineRegion& vgerm() const { return vgerm_; }
  const JunctionRegion& vd_junction() const { return vd_junction_; }
  const GermlineRegion& dgerm() const { return dgerm_; }
  const JunctionRegion& dj_junction() const { return dj_junction_; }
  const GermlineRegion& jgerm() const { return jgerm_; }
  const GermlineRegion& jpadding() const { return jpadding_; }
  const std::vector<double>& vpadding_transition() const { return vpadding_transition_; }
  const MatrixXd& vgerm_vd_junction_transition() const { return vgerm_vd_junction_transition_; }
  const MatrixXd& vd_junction_transition() const { return vd_junction_transition_; }
  const MatrixXd& vd_junction_dgerm_transition() const { return vd_junction_dgerm_transition_; }
  const MatrixXd& dgerm_dj_junction_transition() const { return dgerm_dj_junction_transition_; }
  const MatrixXd& dj_junction_transition() const { return dj_junction_transition_; }
  const MatrixXd& dj_junction_jgerm_transition() const { return dj_junction_jgerm_transition_; }
  const std::vector<double>& jpadding_transition() const { return jpadding_transition_; }
  int vgerm_scaler_count() const { return vgerm_.scaler_count; }
  int dgerm_scaler_count() const { return dgerm_.scaler_count; }
  int jgerm_scaler_count() const { return jgerm_.scaler_count; }
  // sampled annotation of the last SampleNaiveSequence()
  const std::string& vgerm_state_str_samp() const { return vgerm_state_str_samp_; }
  const std::string& dgerm_state_str_samp() const { return dgerm_state_str_samp_; }
  const std::string& jgerm_state_str_samp() const { return jgerm_state_str_samp_; }
  const std::vector<std::string>& vd_junction_state_str_samps() const { return vd_junction_state_str_samps_; }
  const std::vector<std::string>& dj_junction_state_str_samps() const { return dj_junction_state_str_samps_; }
  const std::string& naive_seq_samp() const { return naive_seq_samp_; }
  int vgerm_left_del_samp() const { return vgerm_left_del_samp_; }
  int vgerm_right_del_samp() const { return vgerm_right_del_samp_; }
  int dgerm_left_del_samp() const { return dgerm_left_del_samp_; }
  int dgerm_right_del_samp() const { return dgerm_right_del_samp_; }
  int jgerm_left_del_samp() const { return jgerm_left_del_samp_; }
  int jgerm_right_del_samp() const { return jgerm_right_del_samp_; }
  const std::string& vgerm_left_insertion_samp() const { return vgerm_left_insertion_samp_; }
  const std::string& vd_junction_insertion_samp() const { return vd_junction_insertion_samp_; }
  const std::string& dj_junction_insertion_samp() const { return dj_junction_insertion_samp_; }
  const std::string& jgerm_right_insertion_samp() const { return jgerm_right_insertion_samp_; }

 protected:
  // partis YAML
  std::string locus_;
  lhb::yaml::Node cluster_data_;
  std::map<std::string, std::pair<int, int>> flexbounds_;
  std::map<std::string, int> relpos_;
  std::map<std::string, lhb::GermlineGene> ggenes_;
  std::string alphabet_;  // germline alphabet + "N"
  MatrixXi msa_;          // [sequences x sites]
  std::mt19937 rng_;
  std::discrete_distribution<int> distr_;

  // state space
  GermlineRegion vpadding_, vgerm_, dgerm_, jgerm_, jpadding_;
  JunctionRegion vd_junction_, dj_junction_;
  // transitions
  std::vector<double> vpadding_transition_, jpadding_transition_;
  MatrixXd vgerm_vd_junction_transition_, vd_junction_transition_, vd_junction_dgerm_transition_;
  MatrixXd dgerm_dj_junction_transition_, dj_junction_transition_, dj_junction_jgerm_transition_;

  // sampled path
  std::string naive_seq_samp_;
  std::string vgerm_state_str_samp_, dgerm_state_str_samp_, jgerm_state_str_samp_;
  int vgerm_state_ind_samp_ = -1, dgerm_state_ind_samp_ = -1, jgerm_state_ind_samp_ = -1;
  int vgerm_left_del_samp_ = -1, vgerm_right_del_samp_ = -1, dgerm_left_del_samp_ = -1;
  int dgerm_right_del_samp_ = -1, jgerm_left_del_samp_ = -1, jgerm_right_del_samp_ = -1;
  std::vector<std::string> vd_junction_state_str_samps_, dj_junction_state_str_samps_;
  std::vector<int> vd_junction_state_ind_samps_, dj_junction_state_ind_samps_;
  std::string vd_junction_insertion_samp_, dj_junction_insertion_samp_;
  std::string vgerm_left_insertion_samp_, jgerm_right_insertion_samp_;

  bool cache_forward_ = true;
  // fills the forward tables and scaler counts for the current emissions
  virtual void RunForwardAlgorithm() = 0;

 private:
  void InitializeMsa();
  void InitializeStateSpace();
  void InitializeTransition();
  void SampleInitialState(const int32_t* preset = nullptr);
  std::string SamplePath(const int32_t* idx);
};

class PhyloHMM : public HMM {
 public:
  PhyloHMM(const std::string& yaml_path, int cluster_ind, const std::string& hmm_param_dir, int seed);
  ~PhyloHMM();

  void InitializePhyloParameters(const std::string& newick_path, const std::vector<double>& er,
                                 const std::vector<double>& pi, double alpha, int num_rates);
  void InitializePhyloEmission();
  double LogLikelihood() override;
  // batch_size: RevBayes samples scored per device call (new; the reference scores one at a time). 256: the three
  // stages' working set of a batch (lines, exported trees, forward tables: ~60 KB per row) then stays in the host's
  // caches - measured on 129-tip rows, 16 cores: 77 k rows/s against 65 k with 1024, on a 4096-row trace 57 k / 32 k
  void RunPipeline(const std::string& input_path, const std::string& output_path, int num_rates,
                   int batch_size = 256);

  // Batched evaluation of the loop body src/PhyloHMM.cpp:419-430 without touching the object's
  // current-sample state: log-likelihood of every (newick, er, pi, alpha) sample.
  std::vector<double> ScoreSamples(const std::vector<std::string>& newicks, const std::vector<double>& er,
                                   const std::vector<double>& pi, const std::vector<double>& alpha, int num_rates);
  // pointer form: nothing is copied; Newick ingest of one chunk overlaps the scoring of the previous one
  void ScoreSamples(int n, const char* const* newicks, const double* er, const double* pi, const double* alpha,
                    int num_rates, double* loglik);
  // the same with the strings' lengths (a reader knows them): strings held in pinned memory, one every constant
  // number of bytes, are then copied to the device where they lie (lh_pin_host, include/linearham_b200.h)
  void ScoreSamples(int n, const char* const* newicks, const int64_t* lengths, const double* er, const double* pi,
                    const double* alpha, int num_rates, double* loglik);

  void set_device(int device) { devices_.assign(1, device); }
  // several GPUs of the box: every batched call (ScoreSamples, RunPipeline) cuts its trees into one shard per
  // device (lh_group_t); must be called before the first scoring call
  void set_devices(const std::vector<int>& devices);
  // the device handle behind this object (created on first use); for timings and options
  lh_family* device_family() { EnsureDeviceFamily(); return dev_family_; }
  // host threads for Newick ingest and traversal planning (0 = all hardware threads)
  void set_host_threads(int n);
  // RunPipeline: draw the annotations' states on the device (default) or with the host sampler
  void set_device_sampler(bool on) { device_sampler_ = on; }

  const MatrixXi& xmsa() const { return xmsa_; }
  const std::vector<std::string>& xmsa_labels() const { return xmsa_labels_; }
  const std::vector<std::string>& xmsa_seqs() const { return xmsa_seqs_; }
  int xmsa_naive_ind() const { return xmsa_naive_ind_; }
  const std::vector<double>& xmsa_emission() const { return xmsa_emission_; }
  const std::vector<int>& vpadding_xmsa_inds() const { return vpadding_.xmsa_inds; }
  const std::vector<int>& vgerm_xmsa_inds() const { return vgerm_.xmsa_inds; }
  const MatrixXi& vd_junction_xmsa_inds() const { return vd_junction_.xmsa_inds; }
  const std::vector<int>& dgerm_xmsa_inds() const { return dgerm_.xmsa_inds; }
  const MatrixXi& dj_junction_xmsa_inds() const { return dj_junction_.xmsa_inds; }
  const std::vector<int>& jgerm_xmsa_inds() const { return jgerm_.xmsa_inds; }
  const std::vector<int>& jpadding_xmsa_inds() const { return jpadding_.xmsa_inds; }
  int iteration() const { return iteration_; }
  double rb_loglikelihood() const { return rb_loglikelihood_; }
  double prior() const { return prior_; }
  double alpha() const { return alpha_; }
  const std::vector<double>& er() const { return er_; }
  const std::vector<double>& pi() const { return pi_; }
  const std::string& tree() const { return tree_; }   // Newick of the current sample
  const std::vector<double>& sr() const { return sr_; }
  double lh_loglikelihood() const { return lh_loglikelihood_; }
  double logweight() const { return logweight_; }
  const std::string& naive_sequence() const { return naive_sequence_; }

  // Flat description handed to lh_family_create (pointers stay valid while the object lives).
  struct FlatFamily;
  const lh_family_desc& family_desc() const;

 protected:
  void RunForwardAlgorithm() override;

 private:
  void InitializeXmsaStructs();
  void BuildFlatFamily();
  void EnsureDeviceFamily();
  void ScoreCurrentSample();
  void WriteOutputHeaders(std::ostream& out) const;
  void WriteOutputLine(std::ostream& out) const;

  MatrixXi xmsa_;
  std::vector<std::string> xmsa_labels_, xmsa_seqs_;
  int xmsa_naive_ind_ = 0;
  std::vector<double> xmsa_emission_;
  // current sample
  int iteration_ = 0;
  double rb_loglikelihood_ = 0, prior_ = 0, alpha_ = 1, lh_loglikelihood_ = 0, logweight_ = 0;
  std::vector<double> er_, pi_, sr_;
  std::string tree_, naive_sequence_;
  double loglik_cached_ = 0;
  bool forward_ready_ = false;        // the forward tables of the current sample are in place (ScoreCurrentSample ran)

  std::unique_ptr<FlatFamily> flat_;
  std::string tree_export_cache_;     // RunPipeline: ExportNewick(tree_) made ahead by the reader stage (else empty)
  lh_family* dev_family_ = nullptr;   // the device handle (member 0 of the group when there are several devices)
  lh_group* dev_group_ = nullptr;
  std::vector<int> devices_{0};
  int host_threads_ = 0;
  bool device_sampler_ = true;
};

typedef std::shared_ptr<PhyloHMM> PhyloHMMPtr;

// Product of p[0..n) in the order Eigen's vectorised reduction multiplies a segment starting at
// element `first` of an aligned vector (what `.segment(first, n).prod()` returns in the reference).
double EigenOrderProduct(const double* p, int n, int first);

// pll_utree_export_newick of the parsed tree (utree.c:148-270): the input topology and labels with
// "%f" branch lengths, missing / zero lengths already replaced by EPS, root printed as ":0.0;".
std::string ExportNewick(const std::string& newick);

}  // namespace linearham
