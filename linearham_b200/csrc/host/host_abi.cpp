// Host-side C ABI (include/linearham_b200_host.h). No CUDA in this file.
#include "../../../include/linearham_b200_host.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>

#include "../decimal.hpp"
#include "../plan.hpp"
#include "host_error.hpp"

namespace lhb {
thread_local std::string g_host_err;
int host_fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_host_err = buf;
  return 1;
}
}  // namespace lhb

extern "C" {

const char* lhh_last_error(void) { return lhb::g_host_err.c_str(); }

int lhh_plan_trees(int32_t T, int32_t B, const int32_t* children, const double* blen,
                   const int32_t* root_child, const double* root_blen, int32_t* steps,
                   double* pblen, int32_t* max_slots) {
  static_assert(sizeof(lhb::Step) == 16, "Step must be an int4");
  lhb::PlanScratch sc;
  int mx = 0;
  for (int t = 0; t < B; ++t) {
    int s = lhb::plan_tree(T, children + (size_t)t * (T - 2) * 2, blen + (size_t)t * (T - 2) * 2,
                           root_child[t], root_blen[t],
                           reinterpret_cast<lhb::Step*>(steps) + (size_t)t * (T - 1),
                           pblen + (size_t)t * (2 * T - 3), sc);
    if (s < 0) return lhb::host_fail("tree %d is not a valid post-order operation list", t);
    if (s > mx) mx = s;
  }
  *max_slots = mx;
  return 0;
}

// Test hook for the decimal conversion the device Newick parser uses (decimal.hpp, compiled for the
// host here): n NUL-terminated strings laid end to end in buf.
int lhh_parse_decimal_buf(const char* buf, int32_t n, double* out, int32_t* status, int32_t* consumed) {
  const char* p = buf;
  for (int i = 0; i < n; ++i) {
    const char* q = p;
    out[i] = 0.0;
    status[i] = lhb::parse_decimal([&]() { return (int)(unsigned char)*q; }, [&]() { ++q; }, &out[i]);
    consumed[i] = (int32_t)(q - p);
    p += std::strlen(p) + 1;
  }
  return 0;
}

}  // extern "C"

// ---- C window onto linearham::PhyloHMM (host/phylo_hmm.hpp) ---------------------------------------
#include <algorithm>
#include <map>
#include <stdexcept>
#include <vector>

#include "../../../include/linearham_b200.h"
#include "phylo_hmm.hpp"

namespace {
struct Handle {
  std::unique_ptr<linearham::PhyloHMM> hmm;
};
template <class F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    lhb::g_host_err = e.what();
  } catch (...) {
    lhb::g_host_err = "unknown C++ exception";
  }
  return 1;
}
linearham::PhyloHMM& H(lhh_phylohmm_t* h) { return *reinterpret_cast<Handle*>(h)->hmm; }
}  // namespace

extern "C" {

lhh_phylohmm_t* lhh_phylohmm_new(const char* yaml_path, int32_t cluster_ind, const char* hmm_param_dir, int32_t seed) {
  Handle* h = new Handle;
  if (guarded([&] { h->hmm.reset(new linearham::PhyloHMM(yaml_path, cluster_ind, hmm_param_dir, seed)); })) {
    delete h;
    return nullptr;
  }
  return reinterpret_cast<lhh_phylohmm_t*>(h);
}

void lhh_phylohmm_free(lhh_phylohmm_t* h) { delete reinterpret_cast<Handle*>(h); }

int lhh_phylohmm_set_device(lhh_phylohmm_t* h, int32_t device) {
  return guarded([&] { H(h).set_device(device); });
}
int lhh_phylohmm_set_devices(lhh_phylohmm_t* h, const int32_t* devices, int32_t n_devices) {
  return guarded([&] { H(h).set_devices(std::vector<int>(devices, devices + n_devices)); });
}
int lhh_phylohmm_set_host_threads(lhh_phylohmm_t* h, int32_t n_threads) {
  return guarded([&] { H(h).set_host_threads(n_threads); });
}

int lhh_phylohmm_set_device_sampler(lhh_phylohmm_t* h, int32_t on) {
  return guarded([&] { H(h).set_device_sampler(on != 0); });
}

int lhh_phylohmm_family_desc(lhh_phylohmm_t* h, lh_family_desc_t* out) {
  return guarded([&] { *out = H(h).family_desc(); });
}

int lhh_phylohmm_init_params(lhh_phylohmm_t* h, const char* newick_path, const double* er, const double* pi,
                             double alpha, int32_t num_rates) {
  return guarded([&] {
    H(h).InitializePhyloParameters(newick_path, std::vector<double>(er, er + 6), std::vector<double>(pi, pi + 4), alpha, num_rates);
  });
}

int lhh_phylohmm_init_emission(lhh_phylohmm_t* h) {
  return guarded([&] { H(h).InitializePhyloEmission(); });
}

int lhh_phylohmm_loglikelihood(lhh_phylohmm_t* h, double* out) {
  return guarded([&] { *out = H(h).LogLikelihood(); });
}

int lhh_phylohmm_sample_naive(lhh_phylohmm_t* h, char* out, int32_t cap) {
  return guarded([&] {
    std::string s = H(h).SampleNaiveSequence();
    if ((int)s.size() + 1 > cap) throw std::runtime_error("buffer too small for the naive sequence");
    std::memcpy(out, s.c_str(), s.size() + 1);
  });
}

int lhh_phylohmm_run_pipeline(lhh_phylohmm_t* h, const char* input_path, const char* output_path, int32_t num_rates,
                              int32_t batch_size) {
  return guarded([&] { H(h).RunPipeline(input_path, output_path, num_rates, batch_size); });
}

int lhh_phylohmm_score_samples(lhh_phylohmm_t* h, int32_t n, const char* const* newicks, const double* er, const double* pi,
                               const double* alpha, int32_t num_rates, double* loglik) {
  return guarded([&] {
    H(h).ScoreSamples(n, newicks, er, pi, alpha, num_rates, loglik);
  });
}

int lhh_phylohmm_score_samples_len(lhh_phylohmm_t* h, int32_t n, const char* const* newicks, const int64_t* lengths,
                                   const double* er, const double* pi, const double* alpha, int32_t num_rates,
                                   double* loglik) {
  return guarded([&] {
    H(h).ScoreSamples(n, newicks, lengths, er, pi, alpha, num_rates, loglik);
  });
}

int lhh_phylohmm_device_family(lhh_phylohmm_t* h, lh_family_t** out) {
  return guarded([&] { *out = H(h).device_family(); });
}

int lhh_export_newick(const char* newick, char* out, int32_t cap) {
  return guarded([&] {
    std::string s = linearham::ExportNewick(newick);
    if ((int)s.size() + 1 > cap) throw std::runtime_error("buffer too small for the Newick string");
    std::memcpy(out, s.c_str(), s.size() + 1);
  });
}

// Named read-outs, column-major; pass out = NULL to query the shape.
int lhh_phylohmm_get_f64(lhh_phylohmm_t* h, const char* name, double* out, int32_t* rows, int32_t* cols) {
  return guarded([&] {
    const linearham::PhyloHMM& p = H(h);
    const std::string n(name);
    auto vec = [&](const std::vector<double>& v) {
      *rows = 1;
      *cols = (int)v.size();
      if (out) std::copy(v.begin(), v.end(), out);
    };
    auto mat = [&](const linearham::MatrixXd& m) {
      *rows = m.rows();
      *cols = m.cols();
      if (out) std::copy(m.v.begin(), m.v.end(), out);
    };
    if (n == "xmsa_emission") return vec(p.xmsa_emission());
    if (n == "vpadding_transition") return vec(p.vpadding_transition());
    if (n == "jpadding_transition") return vec(p.jpadding_transition());
    if (n == "vgerm_vd_junction_transition") return mat(p.vgerm_vd_junction_transition());
    if (n == "vd_junction_transition") return mat(p.vd_junction_transition());
    if (n == "vd_junction_dgerm_transition") return mat(p.vd_junction_dgerm_transition());
    if (n == "dgerm_dj_junction_transition") return mat(p.dgerm_dj_junction_transition());
    if (n == "dj_junction_transition") return mat(p.dj_junction_transition());
    if (n == "dj_junction_jgerm_transition") return mat(p.dj_junction_jgerm_transition());
    if (n == "vgerm_forward") return vec(p.vgerm().forward);
    if (n == "dgerm_forward") return vec(p.dgerm().forward);
    if (n == "jgerm_forward") return vec(p.jgerm().forward);
    if (n == "vd_junction_forward") return mat(p.vd_junction().forward);
    if (n == "dj_junction_forward") return mat(p.dj_junction().forward);
    if (n == "sr") return vec(p.sr());
    if (n == "er") return vec(p.er());
    if (n == "pi") return vec(p.pi());
    if (n == "vgerm_gene_prob" || n == "vgerm_trans_prod") {
      const lh_family_desc_t& d = p.family_desc();
      *rows = 1;
      *cols = d.n_v;
      if (out) std::copy_n(n == "vgerm_gene_prob" ? d.v_gene_prob : d.v_trans_prod, d.n_v, out);
      return;
    }
    throw std::runtime_error("lhh_phylohmm_get_f64: unknown field " + n);
  });
}

int lhh_phylohmm_get_i32(lhh_phylohmm_t* h, const char* name, int32_t* out, int32_t* rows, int32_t* cols) {
  return guarded([&] {
    const linearham::PhyloHMM& p = H(h);
    const std::string n(name);
    auto vec = [&](const std::vector<int>& v) {
      *rows = 1;
      *cols = (int)v.size();
      if (out) std::copy(v.begin(), v.end(), out);
    };
    auto mat = [&](const linearham::MatrixXi& m) {
      *rows = m.rows();
      *cols = m.cols();
      if (out) std::copy(m.v.begin(), m.v.end(), out);
    };
    auto one = [&](int v) {
      *rows = *cols = 1;
      if (out) out[0] = v;
    };
    auto ranges = [&](const linearham::GeneRanges& m) {
      *rows = 2;
      *cols = (int)m.size();
      if (out) {
        int i = 0;
        for (const auto& kv : m) {
          out[2 * i] = kv.second.first;
          out[2 * i + 1] = kv.second.second;
          ++i;
        }
      }
    };
    if (n == "msa") return mat(p.msa());
    if (n == "xmsa") return mat(p.xmsa());
    if (n == "xmsa_naive_ind") return one(p.xmsa_naive_ind());
    if (n == "vd_junction_xmsa_inds") return mat(p.vd_junction_xmsa_inds());
    if (n == "dj_junction_xmsa_inds") return mat(p.dj_junction_xmsa_inds());
    if (n == "vgerm_scaler_count") return one(p.vgerm_scaler_count());
    if (n == "dgerm_scaler_count") return one(p.dgerm_scaler_count());
    if (n == "jgerm_scaler_count") return one(p.jgerm_scaler_count());
    if (n == "vd_junction_scaler_counts") return vec(p.vd_junction().scaler_counts);
    if (n == "dj_junction_scaler_counts") return vec(p.dj_junction().scaler_counts);
    const std::map<std::string, const linearham::GermlineRegion*> germ = {
        {"vpadding", &p.vpadding()}, {"vgerm", &p.vgerm()}, {"dgerm", &p.dgerm()}, {"jgerm", &p.jgerm()}, {"jpadding", &p.jpadding()}};
    for (const auto& kv : germ) {
      const std::string& r = kv.first;
      if (n == r + "_xmsa_inds") return vec(kv.second->xmsa_inds);
      if (n == r + "_naive_bases") return vec(kv.second->naive_bases);
      if (n == r + "_germ_inds") return vec(kv.second->germ_inds);
      if (n == r + "_site_inds") return vec(kv.second->site_inds);
      if (n == r + "_left_del") return vec(kv.second->left_del);
      if (n == r + "_right_del") return vec(kv.second->right_del);
      if (n == r + "_ggene_ranges") return ranges(kv.second->ggene_ranges);
    }
    const std::map<std::string, const linearham::JunctionRegion*> junc = {{"vd_junction", &p.vd_junction()}, {"dj_junction", &p.dj_junction()}};
    for (const auto& kv : junc) {
      const std::string& r = kv.first;
      if (n == r + "_naive_bases") return vec(kv.second->naive_bases);
      if (n == r + "_germ_inds") return vec(kv.second->germ_inds);
      if (n == r + "_site_inds") return vec(kv.second->site_inds);
      if (n == r + "_del") return vec(kv.second->del);
      if (n == r + "_ggene_ranges") return ranges(kv.second->ggene_ranges);
    }
    throw std::runtime_error("lhh_phylohmm_get_i32: unknown field " + n);
  });
}

// '\n'-joined strings; returns the size needed (including the NUL) or -1.
int lhh_phylohmm_get_str(lhh_phylohmm_t* h, const char* name, char* out, int32_t cap) {
  std::string s;
  int rc = guarded([&] {
    const linearham::PhyloHMM& p = H(h);
    const std::string n(name);
    auto join = [&](const std::vector<std::string>& v) {
      for (size_t i = 0; i < v.size(); ++i) {
        if (i) s += '\n';
        s += v[i];
      }
    };
    if (n == "xmsa_seqs") join(p.xmsa_seqs());
    else if (n == "xmsa_labels") join(p.xmsa_labels());
    else if (n == "locus") s = p.locus();
    else if (n == "alphabet") s = p.alphabet();
    else if (n == "vgerm_state_strs") join(p.vgerm().state_strs);
    else if (n == "dgerm_state_strs") join(p.dgerm().state_strs);
    else if (n == "jgerm_state_strs") join(p.jgerm().state_strs);
    else if (n == "vd_junction_state_strs") join(p.vd_junction().state_strs);
    else if (n == "dj_junction_state_strs") join(p.dj_junction().state_strs);
    else if (n == "naive_seq_samp") s = p.naive_seq_samp();
    else if (n == "tree") s = p.tree();
    else throw std::runtime_error("lhh_phylohmm_get_str: unknown field " + n);
  });
  if (rc) return -1;
  if (out && cap > 0) {
    std::strncpy(out, s.c_str(), cap - 1);
    out[cap - 1] = 0;
  }
  return (int)s.size() + 1;
}

}  // extern "C"
