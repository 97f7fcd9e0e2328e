// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// extern "C" window onto the UNMODIFIED reference classes
// (linearham::PhyloHMM, /root/reference/src/PhyloHMM.hpp:24-127) so that the
// Python tests / golden-vector generator / cpu_baseline leg of bench.py can
// drive the reference through ctypes. This file is ours; every numeric result
// it returns is produced by reference code compiled from where it lies.
//
// The reference keeps tree_ private and only accepts a Newick *file*
// (InitializePhyloParameters, src/PhyloHMM.cpp:350-361) or a whole TSV
// (RunPipeline, :393-446). To score one Newick *string* per call - what the
// body of the RunPipeline loop does (:419-430) - the shim opens the class up
// with the usual test-only access override.
#include <cstring>
#include <chrono>
#include <fstream>
#include <map>
#include <memory>
#include <random>
#include <regex>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>
#include <Eigen/Dense>
#include <yaml-cpp/yaml.h>

#define private public
#define protected public
#include "PhyloHMM.hpp"
#undef private
#undef protected
#include <pll_util.hpp>
#include "utils.hpp"

using linearham::PhyloHMM;

namespace {
thread_local std::string g_err;

struct Ref {
  std::unique_ptr<PhyloHMM> hmm;
};

template <class F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
  } catch (...) {
    g_err = "unknown C++ exception";
  }
  return 1;
}

// copy a column-major Eigen object out as column-major doubles/ints
template <class M, class T>
void copy_out(const M& m, T* out, int* rows, int* cols) {
  *rows = (int)m.rows();
  *cols = (int)m.cols();
  if (out)
    for (int j = 0; j < m.cols(); ++j)
      for (int i = 0; i < m.rows(); ++i) out[(size_t)j * m.rows() + i] = (T)m(i, j);
}
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

void* ref_phylohmm_new(const char* yaml_path, int cluster_ind,
                       const char* hmm_param_dir, int seed) {
  Ref* r = new Ref;
  int rc = guarded([&] {
    r->hmm.reset(new PhyloHMM(yaml_path, cluster_ind, hmm_param_dir, seed));
    r->hmm->tree_ = nullptr;  // left uninitialised by the reference ctor
  });
  if (rc) {
    delete r;
    return nullptr;
  }
  return r;
}

void ref_phylohmm_free(void* h) {
  Ref* r = (Ref*)h;
  if (!r) return;
  delete r;
}

// One iteration of the RunPipeline loop body up to LogLikelihood()
// (src/PhyloHMM.cpp:419-430), for a Newick string.
int ref_score_newick(void* h, const char* newick, const double* er,
                     const double* pi, double alpha, int num_rates,
                     double* loglik) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  return guarded([&] {
    p.DestroyTree();
    std::string s = std::regex_replace(std::string(newick),
                                       std::regex("\\[\\&index=[0-9]+\\]"), "");
    p.tree_ = pll_utree_parse_newick_string(s.c_str());
    if (!p.tree_) throw std::runtime_error(std::string("newick: ") + pll_errmsg);
    pt::pll::set_missing_branch_length(p.tree_, linearham::EPS);
    p.er_.assign(er, er + 6);
    p.pi_.assign(pi, pi + 4);
    p.alpha_ = alpha;
    p.sr_.assign(num_rates, 0.0);
    pll_compute_gamma_cats(p.alpha_, p.sr_.size(), p.sr_.data(),
                           PLL_GAMMA_RATES_MEAN);
    p.InitializePhyloEmission();
    *loglik = p.LogLikelihood();
  });
}

// Same, timed: scores `n` samples back to back and returns seconds spent in
// InitializePhyloEmission() and LogLikelihood() separately.
int ref_score_batch_timed(void* h, int n, const char* const* newicks,
                          const double* er, const double* pi,
                          const double* alpha, int num_rates, double* loglik,
                          double* sec_parse, double* sec_emission,
                          double* sec_forward) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  using clk = std::chrono::steady_clock;
  double tp = 0, te = 0, tf = 0;
  int rc = guarded([&] {
    std::regex idx("\\[\\&index=[0-9]+\\]");
    for (int i = 0; i < n; ++i) {
      auto t0 = clk::now();
      p.DestroyTree();
      std::string s = std::regex_replace(std::string(newicks[i]), idx, "");
      p.tree_ = pll_utree_parse_newick_string(s.c_str());
      if (!p.tree_) throw std::runtime_error(std::string("newick: ") + pll_errmsg);
      pt::pll::set_missing_branch_length(p.tree_, linearham::EPS);
      p.er_.assign(er + 6 * i, er + 6 * i + 6);
      p.pi_.assign(pi + 4 * i, pi + 4 * i + 4);
      p.alpha_ = alpha[i];
      p.sr_.assign(num_rates, 0.0);
      pll_compute_gamma_cats(p.alpha_, p.sr_.size(), p.sr_.data(),
                             PLL_GAMMA_RATES_MEAN);
      auto t1 = clk::now();
      p.InitializePhyloEmission();
      auto t2 = clk::now();
      loglik[i] = p.LogLikelihood();
      auto t3 = clk::now();
      tp += std::chrono::duration<double>(t1 - t0).count();
      te += std::chrono::duration<double>(t2 - t1).count();
      tf += std::chrono::duration<double>(t3 - t2).count();
    }
  });
  *sec_parse = tp;
  *sec_emission = te;
  *sec_forward = tf;
  return rc;
}

int ref_init_params_file(void* h, const char* newick_path, const double* er,
                         const double* pi, double alpha, int num_rates) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  return guarded([&] {
    p.DestroyTree();
    p.InitializePhyloParameters(newick_path, std::vector<double>(er, er + 6),
                                std::vector<double>(pi, pi + 4), alpha, num_rates);
    p.InitializePhyloEmission();
  });
}

int ref_loglikelihood(void* h, double* out) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  return guarded([&] { *out = p.LogLikelihood(); });
}

int ref_sample_naive(void* h, char* out, int cap) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  return guarded([&] {
    std::string s = p.SampleNaiveSequence();
    std::strncpy(out, s.c_str(), cap - 1);
    out[cap - 1] = 0;
  });
}

int ref_run_pipeline(void* h, const char* in, const char* out, int num_rates) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  return guarded([&] { p.RunPipeline(in, out, num_rates); });
}

int ref_gamma_cats(double alpha, int n, double* out) {
  return pll_compute_gamma_cats(alpha, n, out, PLL_GAMMA_RATES_MEAN) == PLL_SUCCESS ? 0 : 1;
}

// ---- named read-outs ------------------------------------------------------
// Doubles: column-major. Pass out=NULL to query the shape first.
int ref_get_f64(void* h, const char* name, double* out, int* rows, int* cols) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  std::string n(name);
  auto vec = [&](const std::vector<double>& v) {
    *rows = 1; *cols = (int)v.size();
    if (out) std::copy(v.begin(), v.end(), out);
  };
#define LH_M(field) if (n == #field) { copy_out(p.field##_, out, rows, cols); return 0; }
  LH_M(xmsa_emission)
  LH_M(vpadding_transition) LH_M(vgerm_vd_junction_transition) LH_M(vd_junction_transition)
  LH_M(vd_junction_dgerm_transition) LH_M(dgerm_dj_junction_transition)
  LH_M(dj_junction_transition) LH_M(dj_junction_jgerm_transition) LH_M(jpadding_transition)
  LH_M(vpadding_emission) LH_M(vgerm_emission) LH_M(vd_junction_emission) LH_M(dgerm_emission)
  LH_M(dj_junction_emission) LH_M(jgerm_emission) LH_M(jpadding_emission)
  LH_M(vgerm_forward) LH_M(vd_junction_forward) LH_M(dgerm_forward) LH_M(dj_junction_forward)
  LH_M(jgerm_forward)
#undef LH_M
  if (n == "sr") { vec(p.sr_); return 0; }
  if (n == "er") { vec(p.er_); return 0; }
  if (n == "pi") { vec(p.pi_); return 0; }
  // per-V-gene start factors used by ComputeInitialForwardProbabilities
  // (src/HMM.cpp:304-313): gene_prob and prod(transition segment)
  if (n == "vgerm_gene_prob" || n == "vgerm_trans_prod") {
    *rows = 1; *cols = (int)p.vgerm_ggene_ranges_.size();
    if (out) {
      int i = 0;
      for (auto& kv : p.vgerm_ggene_ranges_) {
        const auto& g = p.ggenes_.at(kv.first);
        int s = p.vgerm_germ_inds_[kv.second.first];
        int len = kv.second.second - kv.second.first - 1;
        out[i++] = (n == "vgerm_gene_prob") ? g.germ_ptr->gene_prob()
                                            : g.germ_ptr->transition().segment(s, len).prod();
      }
    }
    return 0;
  }
  g_err = "ref_get_f64: unknown field " + n;
  return 1;
}

int ref_get_i32(void* h, const char* name, int* out, int* rows, int* cols) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  std::string n(name);
  auto vec = [&](const std::vector<int>& v) {
    *rows = 1; *cols = (int)v.size();
    if (out) std::copy(v.begin(), v.end(), out);
  };
  auto one = [&](int v) { *rows = 1; *cols = 1; if (out) out[0] = v; };
#define LH_M(field) if (n == #field) { copy_out(p.field##_, out, rows, cols); return 0; }
#define LH_V(field) if (n == #field) { vec(p.field##_); return 0; }
#define LH_S(field) if (n == #field) { one(p.field##_); return 0; }
  LH_M(msa) LH_M(xmsa)
  LH_M(vpadding_xmsa_inds) LH_M(vgerm_xmsa_inds) LH_M(vd_junction_xmsa_inds)
  LH_M(dgerm_xmsa_inds) LH_M(dj_junction_xmsa_inds) LH_M(jgerm_xmsa_inds)
  LH_M(jpadding_xmsa_inds)
  LH_V(vpadding_naive_bases) LH_V(vpadding_site_inds)
  LH_V(vgerm_naive_bases) LH_V(vgerm_germ_inds) LH_V(vgerm_site_inds)
  LH_V(vgerm_left_del) LH_V(vgerm_right_del)
  LH_V(vd_junction_naive_bases) LH_V(vd_junction_germ_inds) LH_V(vd_junction_site_inds)
  LH_V(vd_junction_del)
  LH_V(dgerm_naive_bases) LH_V(dgerm_germ_inds) LH_V(dgerm_site_inds)
  LH_V(dgerm_left_del) LH_V(dgerm_right_del)
  LH_V(dj_junction_naive_bases) LH_V(dj_junction_germ_inds) LH_V(dj_junction_site_inds)
  LH_V(dj_junction_del)
  LH_V(jgerm_naive_bases) LH_V(jgerm_germ_inds) LH_V(jgerm_site_inds)
  LH_V(jgerm_left_del) LH_V(jgerm_right_del)
  LH_V(jpadding_naive_bases) LH_V(jpadding_site_inds)
  LH_V(vd_junction_scaler_counts) LH_V(dj_junction_scaler_counts)
  LH_S(vgerm_scaler_count) LH_S(dgerm_scaler_count) LH_S(jgerm_scaler_count)
  LH_S(xmsa_naive_ind)
#undef LH_M
#undef LH_V
#undef LH_S
  auto ranges = [&](const std::map<std::string, std::pair<int, int>>& m) {
    *rows = 2; *cols = (int)m.size();
    if (out) { int i = 0; for (auto& kv : m) { out[2*i] = kv.second.first; out[2*i+1] = kv.second.second; ++i; } }
  };
  if (n == "vpadding_ggene_ranges") { ranges(p.vpadding_ggene_ranges_); return 0; }
  if (n == "vgerm_ggene_ranges") { ranges(p.vgerm_ggene_ranges_); return 0; }
  if (n == "vd_junction_ggene_ranges") { ranges(p.vd_junction_ggene_ranges_); return 0; }
  if (n == "dgerm_ggene_ranges") { ranges(p.dgerm_ggene_ranges_); return 0; }
  if (n == "dj_junction_ggene_ranges") { ranges(p.dj_junction_ggene_ranges_); return 0; }
  if (n == "jgerm_ggene_ranges") { ranges(p.jgerm_ggene_ranges_); return 0; }
  if (n == "jpadding_ggene_ranges") { ranges(p.jpadding_ggene_ranges_); return 0; }
  g_err = "ref_get_i32: unknown field " + n;
  return 1;
}

// Strings: '\n'-joined list. Returns needed size (incl. NUL) or -1.
int ref_get_str(void* h, const char* name, char* out, int cap) {
  PhyloHMM& p = *((Ref*)h)->hmm;
  std::string n(name), s;
  auto join = [&](const std::vector<std::string>& v) {
    for (size_t i = 0; i < v.size(); ++i) { if (i) s += '\n'; s += v[i]; }
  };
  if (n == "xmsa_seqs") join(p.xmsa_seqs_);
  else if (n == "xmsa_labels") join(p.xmsa_labels_);
  else if (n == "locus") s = p.locus_;
  else if (n == "alphabet") s = p.alphabet_;
  else if (n == "vgerm_state_strs") join(p.vgerm_state_strs_);
  else if (n == "dgerm_state_strs") join(p.dgerm_state_strs_);
  else if (n == "jgerm_state_strs") join(p.jgerm_state_strs_);
  else if (n == "vd_junction_state_strs") join(p.vd_junction_state_strs_);
  else if (n == "dj_junction_state_strs") join(p.dj_junction_state_strs_);
  else if (n == "naive_seq_samp") s = p.naive_seq_samp_;
  else { g_err = "ref_get_str: unknown field " + n; return -1; }
  if (out && cap > 0) { std::strncpy(out, s.c_str(), cap - 1); out[cap - 1] = 0; }
  return (int)s.size() + 1;
}

}  // extern "C"
