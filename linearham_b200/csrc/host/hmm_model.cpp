// linearham::HMM - per-family construction on the host (partis cluster YAML + HMM parameter
// directory -> MSA, state space, transition matrices) and backward sampling of a naive sequence.
// Restates src/HMM.cpp:27-246, 466-1089 (construction) and :323-431, 1222-1353 (sampling) on plain
// column-major buffers. Transition matrices and state vectors must equal the reference's exactly
// (test/test.cpp compares them with ==); tests/test_host_model.py checks that against the compiled
// reference.
#include <algorithm>
#include <cassert>
#include <cmath>
#include <regex>
#include <stdexcept>

#include "phylo_hmm.hpp"

namespace linearham {

using lhb::GermlineGene;

// Product of p[0..n) in the order Eigen's vectorised reduction (SSE2 packets of two doubles,
// Redux.h LinearVectorizedTraversal) multiplies a segment that starts at element `first` of a
// 16-byte aligned vector; `.prod()` is what the reference calls (src/HMM.cpp:311-314, 868-872).
double EigenOrderProduct(const double* p, int n, int first) {
  if (n <= 0) return 1.0;
  int as = (first % 2 == 0) ? 0 : 1;
  if (as > n) as = n;
  const int size2 = ((n - as) / 4) * 4, size1 = ((n - as) / 2) * 2;
  const int end2 = as + size2, end1 = as + size1;
  double res;
  if (size1) {
    double a0 = p[as], a1 = p[as + 1];
    if (size1 > 2) {
      double b0 = p[as + 2], b1 = p[as + 3];
      for (int i = as + 4; i < end2; i += 4) {
        a0 *= p[i];
        a1 *= p[i + 1];
        b0 *= p[i + 2];
        b1 *= p[i + 3];
      }
      a0 *= b0;
      a1 *= b1;
      if (end1 > end2) {
        a0 *= p[end2];
        a1 *= p[end2 + 1];
      }
    }
    res = a0 * a1;
    for (int i = 0; i < as; ++i) res *= p[i];
    for (int i = end1; i < n; ++i) res *= p[i];
  } else {
    res = p[0];
    for (int i = 1; i < n; ++i) res *= p[i];
  }
  return res;
}

namespace {

// ---- state caching (src/HMM.cpp:466-619) ----------------------------------------------------

void CacheGermlineStates(const GermlineGene& g, std::pair<int, int> lfb, std::pair<int, int> rfb, int relpos,
                         bool left_end, bool right_end, GermlineRegion& R) {
  const int site_start = left_end ? std::max(relpos, lfb.first) : lfb.second;
  const int site_end = right_end ? std::min(relpos + g.length(), rfb.second) : rfb.first;
  const int range_start = (int)R.naive_bases.size();
  R.ggene_ranges.emplace(g.name, std::make_pair(range_start, range_start + (site_end - site_start)));
  R.state_strs.push_back(g.name);
  R.left_del.push_back(site_start - relpos);
  R.right_del.push_back(relpos + g.length() - site_end);
  for (int i = site_start; i < site_end; ++i) {
    R.naive_bases.push_back(g.bases.at(i - relpos));
    R.germ_inds.push_back(i - relpos);
    R.site_inds.push_back(i);
  }
}

void CacheJunctionStates(const GermlineGene& g, std::pair<int, int> lfb, std::pair<int, int> rfb, int relpos,
                         bool left_end, JunctionRegion& R) {
  const int site_start = left_end ? std::max(relpos, lfb.first) : lfb.first;
  const int site_end = left_end ? rfb.second : std::min(relpos + g.length(), rfb.second);
  const int range_start = (int)R.naive_bases.size();
  int range_end = range_start + (site_end - site_start);
  if (left_end) range_end += (int)g.alphabet.size();
  R.ggene_ranges.emplace(g.name, std::make_pair(range_start, range_end));
  if (left_end) {  // the non-templated-insertion states of this gene come first
    for (std::size_t i = 0; i < g.alphabet.size(); ++i) {
      R.state_strs.push_back(g.name + ":N_" + g.alphabet[i]);
      R.del.push_back(-1);
      R.ggene_types.push_back(g.type);
      R.naive_bases.push_back((int)i);
      R.germ_inds.push_back(-1);
      R.site_inds.push_back(-1);
    }
  }
  for (int i = site_start; i < site_end; ++i) {
    R.state_strs.push_back(g.name + ":" + std::to_string(i - relpos));
    R.del.push_back(left_end ? i - relpos : relpos + g.length() - i - 1);
    R.ggene_types.push_back(g.type);
    R.naive_bases.push_back(g.bases.at(i - relpos));
    R.germ_inds.push_back(i - relpos);
    R.site_inds.push_back(i);
  }
}

void CachePaddingStates(const GermlineGene& g, std::pair<int, int> fb, int relpos, bool left_end, GermlineRegion& R) {
  const int site_start = left_end ? fb.first : std::min(relpos + g.length(), fb.second);
  const int site_end = left_end ? std::max(relpos, fb.first) : fb.second;
  const int range_start = (int)R.naive_bases.size();
  R.ggene_ranges.emplace(g.name, std::make_pair(range_start, range_start + (site_end - site_start)));
  for (int i = site_start; i < site_end; ++i) {
    R.naive_bases.push_back((int)g.alphabet.size());
    R.site_inds.push_back(i);
  }
}

// ---- transitions (src/HMM.cpp:647-1089) -----------------------------------------------------

// A window into a matrix: the reference hands FillTransition either the whole matrix, one row of
// it (germline -> junction) or one column (junction -> germline).
struct View {
  MatrixXd* m;
  int r0, c0;
  double& operator()(int i, int j) const { return (*m)(r0 + i, c0 + j); }
};

struct Span {  // what FillTransition needs to know about one side
  int germ_ind_start, site_ind_start, nti_start, nti_length, germ_start, germ_length;
};

void FillTransition(const GermlineGene& from, const GermlineGene& to, char left_gtype, char right_gtype,
                    const Span& row, const Span& col, const View& T) {
  const int A = (int)from.alphabet.size();
  if (from.name == to.name) {  // within one gene
    if (from.type == right_gtype) {
      if (col.nti_length > 0)  // N -> N
        for (int j = 0; j < col.nti_length; ++j)
          for (int i = 0; i < row.nti_length; ++i) T(row.nti_start + i, col.nti_start + j) = from.nti_transition[(std::size_t)j * A + i];
      if (col.germ_length > 0)  // N -> D or N -> J
        for (int j = 0; j < col.germ_length; ++j)
          for (int i = 0; i < row.nti_length; ++i)
            T(row.nti_start + i, col.germ_start + j) = from.nti_landing_out[(std::size_t)(col.germ_ind_start + j) * A + i];
    }
    if (row.germ_length > 0 && col.germ_length > 0) {  // along the gene
      if (row.germ_ind_start == col.germ_ind_start) {
        const int n = std::min(row.germ_length, col.germ_length - 1);
        for (int i = 0; i < n && i < row.germ_length - 1; ++i)
          T(row.germ_start + i, col.germ_start + i + 1) = from.transition.at(row.germ_ind_start + i);
      } else {
        // last position on the row side continues into the first position on the column side
        T(row.germ_start + row.germ_length - 1, col.germ_start) = from.transition.at(row.germ_ind_start + row.germ_length - 1);
      }
    }
  }
  if (from.type == left_gtype && to.type == right_gtype) {  // across genes
    if (row.germ_length > 0 && col.nti_length > 0)  // V -> N or D -> N
      for (int j = 0; j < col.nti_length; ++j)
        for (int i = 0; i < row.germ_length; ++i)
          T(row.germ_start + i, col.nti_start + j) = (from.landing_out.at(row.germ_ind_start + i) * to.gene_prob) * to.nti_landing_in.at(j);
    if (row.germ_length > 0 && col.germ_length > 0) {  // V -> D or D -> J where the genes abut
      int mr = 0, mc = 0;
      bool found = false;
      for (int s = row.site_ind_start; s < row.site_ind_start + row.germ_length && !found; ++s)
        if (s == col.site_ind_start - 1) {
          mr = s - row.site_ind_start;
          mc = 0;
          found = true;
        }
      for (int s = col.site_ind_start + 1; s < col.site_ind_start + col.germ_length && !found; ++s)
        if (row.site_ind_start == s - 1) {
          mr = 0;
          mc = s - col.site_ind_start;
          found = true;
        }
      if (found) {
        const int len = std::min(row.germ_length - mr, col.germ_length - mc);
        for (int i = 0; i < len; ++i)
          T(row.germ_start + mr + i, col.germ_start + mc + i) =
              (from.landing_out.at(row.germ_ind_start + mr + i) * to.gene_prob) * to.landing_in.at(col.germ_ind_start + mc + i);
      }
    }
  }
}

// the rows / columns a gene occupies inside a junction region
Span junction_span(const JunctionRegion& J, const std::pair<int, int>& range, const GermlineGene& g, char right_gtype) {
  Span s;
  s.nti_start = range.first;
  s.nti_length = (g.type == right_gtype) ? (int)g.alphabet.size() : 0;
  s.germ_start = range.first + s.nti_length;
  s.germ_length = range.second - s.germ_start;
  s.germ_ind_start = s.germ_length > 0 ? J.germ_inds[s.germ_start] : -1;
  s.site_ind_start = s.germ_length > 0 ? J.site_inds[s.germ_start] : -1;
  return s;
}

void ComputeGermlineJunctionTransition(const GermlineRegion& G, const JunctionRegion& J, char left_gtype, char right_gtype,
                                       const std::map<std::string, GermlineGene>& genes, MatrixXd& T) {
  T.assign((int)G.state_strs.size(), (int)J.state_strs.size(), 0.0);
  int from_i = 0;
  for (auto f = G.ggene_ranges.begin(); f != G.ggene_ranges.end(); ++f, ++from_i) {
    const GermlineGene& from = genes.at(f->first);
    // the germline state hands over from its last site
    const Span row{G.germ_inds.at(f->second.second - 1), G.site_inds.at(f->second.second - 1), 0, 0, 0, 1};
    for (auto t = J.ggene_ranges.begin(); t != J.ggene_ranges.end(); ++t) {
      const GermlineGene& to = genes.at(t->first);
      FillTransition(from, to, left_gtype, right_gtype, row, junction_span(J, t->second, to, right_gtype), View{&T, from_i, 0});
    }
  }
}

void ComputeJunctionTransition(const JunctionRegion& J, char left_gtype, char right_gtype,
                               const std::map<std::string, GermlineGene>& genes, MatrixXd& T) {
  T.assign((int)J.state_strs.size(), (int)J.state_strs.size(), 0.0);
  for (auto f = J.ggene_ranges.begin(); f != J.ggene_ranges.end(); ++f) {
    const GermlineGene& from = genes.at(f->first);
    const Span row = junction_span(J, f->second, from, right_gtype);
    for (auto t = J.ggene_ranges.begin(); t != J.ggene_ranges.end(); ++t) {
      const GermlineGene& to = genes.at(t->first);
      FillTransition(from, to, left_gtype, right_gtype, row, junction_span(J, t->second, to, right_gtype), View{&T, 0, 0});
    }
  }
}

void ComputeJunctionGermlineTransition(const JunctionRegion& J, const GermlineRegion& G, char left_gtype, char right_gtype,
                                       const std::map<std::string, GermlineGene>& genes, MatrixXd& T) {
  T.assign((int)J.state_strs.size(), (int)G.state_strs.size(), 0.0);
  for (auto f = J.ggene_ranges.begin(); f != J.ggene_ranges.end(); ++f) {
    const GermlineGene& from = genes.at(f->first);
    const Span row = junction_span(J, f->second, from, right_gtype);
    int to_i = 0;
    for (auto t = G.ggene_ranges.begin(); t != G.ggene_ranges.end(); ++t, ++to_i) {
      const GermlineGene& to = genes.at(t->first);
      // the germline state is entered at its first site
      const Span col{G.germ_inds.at(t->second.first), G.site_inds.at(t->second.first), 0, 0, 0, 1};
      FillTransition(from, to, left_gtype, right_gtype, row, col, View{&T, 0, to_i});
      // ... and then runs through its whole region
      const int len = t->second.second - t->second.first - 1;
      const double through = EigenOrderProduct(to.transition.data() + col.germ_ind_start, len, col.germ_ind_start);
      for (int i = f->second.first; i < f->second.second; ++i) T(i, to_i) *= through;
    }
  }
}

void ComputePaddingTransition(const GermlineRegion& P, const std::map<std::string, GermlineGene>& genes,
                              std::vector<double>& out) {
  out.assign(P.ggene_ranges.size(), 0.0);
  int i = 0;
  for (auto it = P.ggene_ranges.begin(); it != P.ggene_ranges.end(); ++it, ++i) {
    const double n = genes.at(it->first).n_transition;
    out[i] = (1.0 - n) * std::pow(n, it->second.second - it->second.first);
  }
}

// ---- sampling (src/HMM.cpp:1222-1353) -------------------------------------------------------

void SampleJunctionStates(int germ_state_ind, const MatrixXd& junction_germ_T, const JunctionRegion& J,
                          const MatrixXd& junction_T, char left_gtype, char right_gtype, int site_start,
                          const std::string& alphabet, std::mt19937& rng, std::discrete_distribution<int>& distr,
                          std::string& naive, int& germ_left_del, std::vector<std::string>& strs, std::vector<int>& inds,
                          std::string& insertion, int& germ_right_del, const int32_t* preset = nullptr) {
  // preset: the states drawn elsewhere (k_sample, lh_sample_paths), site 0 first - everything but the draws runs here
  const int L = preset ? J.xmsa_inds.rows() : J.forward.rows(), S = preset ? 0 : J.forward.cols();
  strs.assign(L, "");
  inds.assign(L, -1);
  insertion.clear();
  germ_right_del = -1;
  std::vector<double> probs(S);
  for (int i = L - 1; i >= 0; --i) {
    int k;
    if (preset) {
      k = preset[i];
    } else {
      const double* column = (i == L - 1) ? junction_germ_T.data() + (std::size_t)germ_state_ind * junction_germ_T.rows()
                                          : junction_T.data() + (std::size_t)inds[i + 1] * junction_T.rows();
      for (int j = 0; j < S; ++j) probs[j] = column[j] * J.forward(i, j);
      distr.param(std::discrete_distribution<int>::param_type(probs.begin(), probs.end()));
      k = distr(rng);
    }
    inds[i] = k;
    strs[i] = J.state_strs[k];
    naive[site_start + i] = alphabet[J.naive_bases[k]];
    if (J.ggene_types[k] == right_gtype) {
      if (J.del[k] != -1) germ_left_del = J.del[k];
      else insertion = alphabet[J.naive_bases[k]] + insertion;
    } else if (J.ggene_types[k] == left_gtype && germ_right_del == -1) {
      germ_right_del = J.del[k];
    }
  }
}

void SampleGermlineState(const std::vector<int>& junction_inds, const MatrixXd& germ_junction_T, const GermlineRegion& G,
                         const std::string& alphabet, std::mt19937& rng, std::discrete_distribution<int>& distr,
                         std::string& naive, std::string& str, int& ind, int& left_del, int& right_del,
                         const int32_t* preset = nullptr) {
  if (preset) {
    ind = *preset;
  } else {
    const int n = (int)G.state_strs.size();
    std::vector<double> probs(n);
    const double* column = germ_junction_T.data() + (std::size_t)junction_inds.front() * germ_junction_T.rows();
    for (int j = 0; j < n; ++j) probs[j] = column[j] * G.forward[j];
    distr.param(std::discrete_distribution<int>::param_type(probs.begin(), probs.end()));
    ind = distr(rng);
  }
  str = G.state_strs[ind];
  left_del = G.left_del[ind];
  if (right_del == -1) right_del = G.right_del[ind];
  const std::pair<int, int>& range = G.ggene_ranges.at(str);
  for (int i = range.first; i < range.second; ++i) naive[G.site_inds[i]] = alphabet[G.naive_bases[i]];
}

}  // namespace

// ---------------------------------------------------------------------------------------------

HMM::HMM(const std::string& yaml_path, int cluster_ind, const std::string& hmm_param_dir, int seed) {
  lhb::yaml::Node root = lhb::yaml::load_file(yaml_path);
  try {
    locus_ = root.at("germline-info").at("locus").as_string();
    cluster_data_ = root.at("events").at((std::size_t)cluster_ind);
  } catch (const std::runtime_error&) {
    throw std::runtime_error("Can't read one of \"locus\" (under \"germline-info\"), \"events\" from  " + yaml_path +
                             " . Check that yaml file contains these keys.");
  }
  try {
    const lhb::yaml::Node& info = cluster_data_.at("linearham-info");
    for (const auto& kv : info.at("flexbounds").map)
      flexbounds_[kv.first] = std::make_pair(kv.second.at((std::size_t)0).as_int(), kv.second.at((std::size_t)1).as_int());
    for (const auto& kv : info.at("relpos").map) relpos_[kv.first] = kv.second.as_int();
    if (flexbounds_.empty() || relpos_.empty()) throw std::runtime_error("empty");
  } catch (const std::runtime_error&) {
    throw std::runtime_error("Can't read one of \"flexbounds\", \"relpos\" from  " + yaml_path +
                             " . Check that \"linearham-info\" was written to the yaml file and is not null.");
  }
  if (locus_ != "igh" && locus_ != "igk" && locus_ != "igl") throw std::runtime_error("unknown locus \"" + locus_ + "\"");
  ggenes_ = lhb::load_germline_dir(hmm_param_dir);
  alphabet_ = ggenes_.begin()->second.alphabet + "N";
  InitializeMsa();
  rng_.seed(seed);
  InitializeStateSpace();
  InitializeTransition();
}

// src/HMM.cpp:71-83
void HMM::InitializeMsa() {
  const int n = (int)cluster_data_.at("unique_ids").size();
  const int L = (int)cluster_data_.at("naive_seq").as_string().size();
  msa_.assign(n, L, -1);
  for (int i = 0; i < n; ++i) {
    const char* key = cluster_data_.at("has_shm_indels").at((std::size_t)i).as_bool() ? "indel_reversed_seqs" : "input_seqs";
    const std::string& seq = cluster_data_.at(key).at((std::size_t)i).as_string();
    if ((int)seq.size() != L) throw std::runtime_error("sequence " + std::to_string(i) + " does not have the length of naive_seq");
    for (int j = 0; j < L; ++j) {
      const std::size_t b = alphabet_.find(seq[j]);
      if (b == std::string::npos) throw std::runtime_error(std::string("base '") + seq[j] + "' is outside the alphabet " + alphabet_);
      msa_(i, j) = (int)b;
    }
  }
}

// src/HMM.cpp:94-185
void HMM::InitializeStateSpace() {
  const bool igh = locus_ == "igh";
  auto fb = [&](const char* k) {
    auto it = flexbounds_.find(k);
    if (it == flexbounds_.end()) throw std::runtime_error(std::string("flexbounds has no \"") + k + "\"");
    return it->second;
  };
  for (const auto& kv : relpos_) {  // lexicographic gene order, as std::map in the reference
    auto gi = ggenes_.find(kv.first);
    if (gi == ggenes_.end()) throw std::out_of_range("relpos gene " + kv.first + " has no HMM parameter file");
    const GermlineGene& g = gi->second;
    const int relpos = kv.second;
    if (g.type == 'V') {
      CachePaddingStates(g, fb("v_l"), relpos, true, vpadding_);
      CacheGermlineStates(g, fb("v_l"), fb("v_r"), relpos, true, false, vgerm_);
      CacheJunctionStates(g, fb("v_r"), fb(igh ? "d_l" : "j_l"), relpos, false, vd_junction_);
    } else if (g.type == 'D') {
      CacheJunctionStates(g, fb("v_r"), fb("d_l"), relpos, true, vd_junction_);
      CacheGermlineStates(g, fb("d_l"), fb("d_r"), relpos, false, false, dgerm_);
      CacheJunctionStates(g, fb("d_r"), fb("j_l"), relpos, false, dj_junction_);
    } else {
      if (igh) CacheJunctionStates(g, fb("d_r"), fb("j_l"), relpos, true, dj_junction_);
      else CacheJunctionStates(g, fb("v_r"), fb("j_l"), relpos, true, vd_junction_);
      CacheGermlineStates(g, fb("j_l"), fb("j_r"), relpos, false, true, jgerm_);
      CachePaddingStates(g, fb("j_r"), relpos, false, jpadding_);
    }
  }
}

// src/HMM.cpp:190-246
void HMM::InitializeTransition() {
  ComputePaddingTransition(vpadding_, ggenes_, vpadding_transition_);
  if (locus_ == "igh") {
    ComputeGermlineJunctionTransition(vgerm_, vd_junction_, 'V', 'D', ggenes_, vgerm_vd_junction_transition_);
    ComputeJunctionTransition(vd_junction_, 'V', 'D', ggenes_, vd_junction_transition_);
    ComputeJunctionGermlineTransition(vd_junction_, dgerm_, 'V', 'D', ggenes_, vd_junction_dgerm_transition_);
    ComputeGermlineJunctionTransition(dgerm_, dj_junction_, 'D', 'J', ggenes_, dgerm_dj_junction_transition_);
    ComputeJunctionTransition(dj_junction_, 'D', 'J', ggenes_, dj_junction_transition_);
    ComputeJunctionGermlineTransition(dj_junction_, jgerm_, 'D', 'J', ggenes_, dj_junction_jgerm_transition_);
  } else {
    ComputeGermlineJunctionTransition(vgerm_, vd_junction_, 'V', 'J', ggenes_, vgerm_vd_junction_transition_);
    ComputeJunctionTransition(vd_junction_, 'V', 'J', ggenes_, vd_junction_transition_);
    ComputeJunctionGermlineTransition(vd_junction_, jgerm_, 'V', 'J', ggenes_, vd_junction_dgerm_transition_);
  }
  ComputePaddingTransition(jpadding_, ggenes_, jpadding_transition_);
}

// src/HMM.cpp:323-341
void HMM::SampleInitialState(const int32_t* preset) {
  if (preset) {
    jgerm_state_ind_samp_ = *preset;
  } else {
    distr_.param(std::discrete_distribution<int>::param_type(jgerm_.forward.begin(), jgerm_.forward.end()));
    jgerm_state_ind_samp_ = distr_(rng_);
  }
  jgerm_state_str_samp_ = jgerm_.state_strs[jgerm_state_ind_samp_];
  jgerm_left_del_samp_ = jgerm_.left_del[jgerm_state_ind_samp_];
  jgerm_right_del_samp_ = jgerm_.right_del[jgerm_state_ind_samp_];
  const std::pair<int, int>& range = jgerm_.ggene_ranges.at(jgerm_state_str_samp_);
  for (int i = range.first; i < range.second; ++i) naive_seq_samp_[jgerm_.site_inds[i]] = alphabet_[jgerm_.naive_bases[i]];
}

// src/HMM.cpp:358-431
std::string HMM::SampleNaiveSequence() {
  if (cache_forward_) {
    RunForwardAlgorithm();
    cache_forward_ = false;
  }
  return SamplePath(nullptr);
}

// The annotation of a path whose states were drawn elsewhere (k_sample through lh_sample_paths): J gene | D-J states,
// site 0 first | D gene | V-D states | V gene (light chains: J gene | V-J states | V gene). The forward tables are not
// read; everything SampleNaiveSequence derives from the drawn states is derived here by the same code.
std::string HMM::ApplySampledPath(const int32_t* idx) { return SamplePath(idx); }

std::string HMM::SamplePath(const int32_t* idx) {
  const int32_t* jgerm_idx = idx;
  const int32_t *dj_idx = nullptr, *dgerm_idx = nullptr, *vd_idx = nullptr, *vgerm_idx = nullptr;
  if (idx) {
    const int32_t* p = idx + 1;
    if (locus_ == "igh") {
      dj_idx = p;
      p += dj_junction_.xmsa_inds.rows();
      dgerm_idx = p++;
    }
    vd_idx = p;
    p += vd_junction_.xmsa_inds.rows();
    vgerm_idx = p;
  }
  naive_seq_samp_.assign(msa_.cols(), 'N');
  SampleInitialState(jgerm_idx);
  if (locus_ == "igh") {
    SampleJunctionStates(jgerm_state_ind_samp_, dj_junction_jgerm_transition_, dj_junction_, dj_junction_transition_, 'D', 'J',
                         flexbounds_.at("d_r").first, alphabet_, rng_, distr_, naive_seq_samp_, jgerm_left_del_samp_,
                         dj_junction_state_str_samps_, dj_junction_state_ind_samps_, dj_junction_insertion_samp_,
                         dgerm_right_del_samp_, dj_idx);
    SampleGermlineState(dj_junction_state_ind_samps_, dgerm_dj_junction_transition_, dgerm_, alphabet_, rng_, distr_,
                        naive_seq_samp_, dgerm_state_str_samp_, dgerm_state_ind_samp_, dgerm_left_del_samp_,
                        dgerm_right_del_samp_, dgerm_idx);
    SampleJunctionStates(dgerm_state_ind_samp_, vd_junction_dgerm_transition_, vd_junction_, vd_junction_transition_, 'V', 'D',
                         flexbounds_.at("v_r").first, alphabet_, rng_, distr_, naive_seq_samp_, dgerm_left_del_samp_,
                         vd_junction_state_str_samps_, vd_junction_state_ind_samps_, vd_junction_insertion_samp_,
                         vgerm_right_del_samp_, vd_idx);
  } else {
    SampleJunctionStates(jgerm_state_ind_samp_, vd_junction_dgerm_transition_, vd_junction_, vd_junction_transition_, 'V', 'J',
                         flexbounds_.at("v_r").first, alphabet_, rng_, distr_, naive_seq_samp_, jgerm_left_del_samp_,
                         vd_junction_state_str_samps_, vd_junction_state_ind_samps_, vd_junction_insertion_samp_,
                         vgerm_right_del_samp_, vd_idx);
  }
  SampleGermlineState(vd_junction_state_ind_samps_, vgerm_vd_junction_transition_, vgerm_, alphabet_, rng_, distr_,
                      naive_seq_samp_, vgerm_state_str_samp_, vgerm_state_ind_samp_, vgerm_left_del_samp_,
                      vgerm_right_del_samp_, vgerm_idx);
  // V / J framework insertions: the leading and trailing runs of N around a run of alphabet
  // bases (GetFrameworkInsertionRegex, src/utils.cpp:97-99: ^(N*)[ACGT]+(N*)$)
  vgerm_left_insertion_samp_.clear();
  jgerm_right_insertion_samp_.clear();
  std::size_t a = 0, b = naive_seq_samp_.size();
  while (a < b && naive_seq_samp_[a] == 'N') ++a;
  while (b > a && naive_seq_samp_[b - 1] == 'N') --b;
  const std::string germ_alphabet = alphabet_.substr(0, alphabet_.size() - 1);
  if (b > a && naive_seq_samp_.find_first_not_of(germ_alphabet, a) >= b) {
    vgerm_left_insertion_samp_ = naive_seq_samp_.substr(0, a);
    jgerm_right_insertion_samp_ = naive_seq_samp_.substr(b);
  }
  return naive_seq_samp_;
}

}  // namespace linearham
