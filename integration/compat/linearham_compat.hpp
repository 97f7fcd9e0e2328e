// Source-compatibility adapter: the flat, Eigen-typed accessor surface of the reference's HMM / PhyloHMM classes
// (src/HMM.hpp:197-409, src/PhyloHMM.hpp:60-127) over the product's classes (linearham_b200/csrc/host/phylo_hmm.hpp),
// which group the same members by region (vgerm().state_strs) and hold matrices as plain column-major buffers.
// Code written against the reference - its own Catch file first of all, test/test.cpp:750-1399 - compiles against
// this header unchanged; integration/compat/Makefile builds exactly that TEST_CASE from where it lies in the
// reference tree and tests/test_gpu_catch_compat.py runs it on the GPU.
//
// Needs Eigen (the reference tree's lib/eigen; the product itself does not use it), so it lives here and not in
// the product's include directory. Accessors return by value where the product's storage is not the Eigen type.
// Not provided (no counterpart in the product: the data never exists on the host, or is an implementation detail of
// the reference): cluster_data() (YAML::Node), ggenes() as std::unordered_map, rng(), distr(), the per-region
// *_emission() tables (the emission fill happens on the device, only xmsa_emission() comes back), partition()
// and tree() as libpll objects (tree() is the Newick string in the product).
#pragma once
#include <Eigen/Dense>

#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "host/phylo_hmm.hpp"

namespace linearham_compat {

enum class GermlineType { V, D, J };   // src/VDJGermline.hpp:45

typedef std::map<std::string, std::pair<int, int>> GeneRanges;

inline Eigen::MatrixXd to_eigen(const linearham::MatrixXd& m) {
  return Eigen::Map<const Eigen::MatrixXd>(m.data(), m.rows(), m.cols());
}
inline Eigen::MatrixXi to_eigen(const linearham::MatrixXi& m) {
  return Eigen::Map<const Eigen::MatrixXi>(m.data(), m.rows(), m.cols());
}
inline Eigen::RowVectorXd to_row(const std::vector<double>& v) {
  return Eigen::Map<const Eigen::RowVectorXd>(v.data(), (Eigen::Index)v.size());
}
inline Eigen::VectorXd to_col(const std::vector<double>& v) {
  return Eigen::Map<const Eigen::VectorXd>(v.data(), (Eigen::Index)v.size());
}
inline Eigen::VectorXi to_col(const std::vector<int>& v) {
  return Eigen::Map<const Eigen::VectorXi>(v.data(), (Eigen::Index)v.size());
}
inline std::vector<GermlineType> to_types(const std::vector<char>& v) {
  std::vector<GermlineType> out;
  out.reserve(v.size());
  for (char c : v) out.push_back(c == 'V' ? GermlineType::V : c == 'D' ? GermlineType::D : GermlineType::J);
  return out;
}

// One line per reference accessor: name, return type, where the product keeps it.
#define LHC_GERMLINE_REGION(r)                                                                        \
  const std::vector<std::string>& r##_state_strs() const { return r().state_strs; }                   \
  const std::vector<int>& r##_left_del() const { return r().left_del; }                               \
  const std::vector<int>& r##_right_del() const { return r().right_del; }                             \
  const GeneRanges& r##_ggene_ranges() const { return r().ggene_ranges; }                             \
  const std::vector<int>& r##_naive_bases() const { return r().naive_bases; }                         \
  const std::vector<int>& r##_germ_inds() const { return r().germ_inds; }                             \
  const std::vector<int>& r##_site_inds() const { return r().site_inds; }                             \
  Eigen::VectorXi r##_xmsa_inds() const { return to_col(r().xmsa_inds); }                             \
  Eigen::RowVectorXd r##_forward() const { return to_row(r().forward); }

#define LHC_PADDING_REGION(r)                                                                         \
  const GeneRanges& r##_ggene_ranges() const { return r().ggene_ranges; }                             \
  const std::vector<int>& r##_naive_bases() const { return r().naive_bases; }                         \
  const std::vector<int>& r##_site_inds() const { return r().site_inds; }                             \
  Eigen::VectorXi r##_xmsa_inds() const { return to_col(r().xmsa_inds); }

#define LHC_JUNCTION_REGION(r)                                                                        \
  const std::vector<std::string>& r##_state_strs() const { return r().state_strs; }                   \
  const std::vector<int>& r##_del() const { return r().del; }                                         \
  std::vector<GermlineType> r##_ggene_types() const { return to_types(r().ggene_types); }             \
  const GeneRanges& r##_ggene_ranges() const { return r().ggene_ranges; }                             \
  const std::vector<int>& r##_naive_bases() const { return r().naive_bases; }                         \
  const std::vector<int>& r##_germ_inds() const { return r().germ_inds; }                             \
  const std::vector<int>& r##_site_inds() const { return r().site_inds; }                             \
  Eigen::MatrixXi r##_xmsa_inds() const { return to_eigen(r().xmsa_inds); }                           \
  Eigen::MatrixXd r##_forward() const { return to_eigen(r().forward); }                               \
  const std::vector<int>& r##_scaler_counts() const { return r().scaler_counts; }

class PhyloHMM : public linearham::PhyloHMM {
  typedef linearham::PhyloHMM Base;

 public:
  using Base::Base;   // (yaml_path, cluster_ind, hmm_param_dir, seed), src/PhyloHMM.hpp:57-58

  Eigen::MatrixXi msa() const { return to_eigen(Base::msa()); }
  LHC_PADDING_REGION(vpadding)
  LHC_GERMLINE_REGION(vgerm)
  LHC_JUNCTION_REGION(vd_junction)
  LHC_GERMLINE_REGION(dgerm)
  LHC_JUNCTION_REGION(dj_junction)
  LHC_GERMLINE_REGION(jgerm)
  LHC_PADDING_REGION(jpadding)

  Eigen::RowVectorXd vpadding_transition() const { return to_row(Base::vpadding_transition()); }
  Eigen::MatrixXd vgerm_vd_junction_transition() const { return to_eigen(Base::vgerm_vd_junction_transition()); }
  Eigen::MatrixXd vd_junction_transition() const { return to_eigen(Base::vd_junction_transition()); }
  Eigen::MatrixXd vd_junction_dgerm_transition() const { return to_eigen(Base::vd_junction_dgerm_transition()); }
  Eigen::MatrixXd dgerm_dj_junction_transition() const { return to_eigen(Base::dgerm_dj_junction_transition()); }
  Eigen::MatrixXd dj_junction_transition() const { return to_eigen(Base::dj_junction_transition()); }
  Eigen::MatrixXd dj_junction_jgerm_transition() const { return to_eigen(Base::dj_junction_jgerm_transition()); }
  Eigen::RowVectorXd jpadding_transition() const { return to_row(Base::jpadding_transition()); }

  bool cache_forward() const { return cache_forward_; }
  int vgerm_state_ind_samp() const { return vgerm_state_ind_samp_; }
  int dgerm_state_ind_samp() const { return dgerm_state_ind_samp_; }
  int jgerm_state_ind_samp() const { return jgerm_state_ind_samp_; }
  const std::vector<int>& vd_junction_state_ind_samps() const { return vd_junction_state_ind_samps_; }
  const std::vector<int>& dj_junction_state_ind_samps() const { return dj_junction_state_ind_samps_; }

  Eigen::MatrixXi xmsa() const { return to_eigen(Base::xmsa()); }
  Eigen::VectorXd xmsa_emission() const { return to_col(Base::xmsa_emission()); }
};

typedef std::shared_ptr<PhyloHMM> PhyloHMMPtr;

#undef LHC_GERMLINE_REGION
#undef LHC_PADDING_REGION
#undef LHC_JUNCTION_REGION

}  // namespace linearham_compat
