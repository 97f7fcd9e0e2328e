// linearham::PhyloHMM - xMSA construction (bit-exact restatement of src/PhyloHMM.cpp:45-89,
// 123-144, 461-536), the per-sample entry points of the reference API routed to the CUDA library,
// and the batched RunPipeline (src/PhyloHMM.cpp:393-446 with B rows per device call).
#include "phylo_hmm.hpp"

#include <charconv>
#include <chrono>
#include <cstdlib>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <future>
#include <sstream>
#include <stdexcept>

#include "../pool.hpp"
#include "../../../include/linearham_b200.h"
#include "../../../include/linearham_b200_host.h"

namespace lhb {
int gamma_mean_rates(double alpha, int n, double* out);  // gamma.cpp
}

namespace linearham {

namespace {

const double kEps = 1e-6;  // src/utils.hpp:20

// ({naive base, MSA site} -> xMSA column), columns numbered in first-insertion order
typedef std::map<std::pair<int, int>, int> XmsaIds;

int StoreXmsaIndex(std::pair<int, int> id, XmsaIds& ids) {
  auto ins = ids.emplace(id, (int)ids.size());
  return ins.first->second;
}

void StoreGermlinePaddingXmsaIndices(GermlineRegion& R, XmsaIds& ids) {
  R.xmsa_inds.assign(R.naive_bases.size(), -1);
  for (std::size_t i = 0; i < R.naive_bases.size(); ++i) R.xmsa_inds[i] = StoreXmsaIndex({R.naive_bases[i], R.site_inds[i]}, ids);
}

void StoreJunctionXmsaIndices(JunctionRegion& R, std::pair<int, int> lfb, std::pair<int, int> rfb, XmsaIds& ids) {
  const int site_start = lfb.first, site_end = rfb.second;
  R.xmsa_inds.assign(site_end - site_start, (int)R.naive_bases.size(), -1);
  for (std::size_t i = 0; i < R.naive_bases.size(); ++i) {
    if (R.site_inds[i] == -1) {  // an NTI state may sit at every junction site
      for (int s = site_start; s < site_end; ++s) R.xmsa_inds(s - site_start, (int)i) = StoreXmsaIndex({R.naive_bases[i], s}, ids);
    } else {
      R.xmsa_inds(R.site_inds[i] - site_start, (int)i) = StoreXmsaIndex({R.naive_bases[i], R.site_inds[i]}, ids);
    }
  }
}

std::string read_file(const std::string& path) {
  std::ifstream in(path);
  if (!in) throw std::runtime_error("cannot open " + path);
  std::stringstream ss;
  ss << in.rdbuf();
  return ss.str();
}

// one line of a tab-separated file with "" quoting (fast-cpp-csv-parser double_quote_escape<'\t','"'>)
std::vector<std::string> split_tsv(const std::string& line) {
  std::vector<std::string> out;
  // (a row without a quote or a carriage return - every row RevBayes writes - is cut at its tabs with memchr: the
  // character loop below costs as much per 9 KB row as re-exporting its tree)
  if (!std::memchr(line.data(), '"', line.size()) && !std::memchr(line.data(), '\r', line.size())) {
    const char* p = line.data();
    const char* const end = p + line.size();
    for (;;) {
      const char* tab = static_cast<const char*>(std::memchr(p, '\t', (std::size_t)(end - p)));
      out.emplace_back(p, tab ? tab : end);
      if (!tab) return out;
      p = tab + 1;
    }
  }
  std::string cur;
  bool quoted = false;
  for (std::size_t i = 0; i < line.size(); ++i) {
    const char c = line[i];
    if (quoted) {
      if (c == '"') {
        if (i + 1 < line.size() && line[i + 1] == '"') { cur.push_back('"'); ++i; }
        else quoted = false;
      } else {
        cur.push_back(c);
      }
    } else if (c == '"') {
      quoted = true;
    } else if (c == '\t') {
      out.push_back(cur);
      cur.clear();
    } else if (c != '\r') {
      cur.push_back(c);
    }
  }
  out.push_back(cur);
  return out;
}

double to_double(const std::string& s, const std::string& what) {
  char* end = nullptr;
  const double v = std::strtod(s.c_str(), &end);
  if (end == s.c_str()) throw std::runtime_error("cannot parse " + what + " \"" + s + "\"");
  return v;
}

// ---- Newick re-export (pll_utree_export_newick, utree.c:148-270) ----------------------------
// The exported text lists the nodes in the order the input does, so the export is one streaming pass over the
// characters: comments and blanks go, every length becomes "%f" of the parsed value (a missing or zero length EPS,
// src/PhyloHMM.cpp:422), the trifurcation ends in ":0.0;". std::to_chars(fixed, 6) prints the digits printf's "%f"
// prints (both round the exact binary value). A 9 KB string took 218 us when the text was assembled subtree by
// subtree; this pass takes a tenth of that, which matters because RunPipeline's reader does it for every row.
inline bool nw_delim(char c) {
  return c == '(' || c == ')' || c == ',' || c == ':' || c == ';' || c == '[' || c == ' ' || c == '\t' || c == '\n' ||
         c == '\r' || c == '\0';
}
inline const char* nw_skip(const char* p) {
  for (;;) {
    while (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r') ++p;
    if (*p != '[') return p;
    while (*p != ']' && *p) ++p;
    if (*p) ++p;
  }
}
// optional ":number" -> its value (0 when absent); the reference reads it with atof
inline const char* nw_length(const char* p, const char* end, double* out) {
  p = nw_skip(p);
  *out = 0.0;
  if (*p != ':') return p;
  p = nw_skip(p + 1);
  if ((*p >= '0' && *p <= '9') || *p == '-' || *p == '.') {
    const auto r = std::from_chars(p, end, *out);
    if (r.ec == std::errc() && nw_delim(*r.ptr)) return r.ptr;
  }
  char* stop = nullptr;
  *out = std::strtod(p, &stop);
  if (stop == p) throw std::invalid_argument("malformed branch length in Newick string");
  return stop;
}
inline void nw_put_length(std::string& out, double v) {
  char buf[512];  // "%f" of 1e308 has 316 characters
  out.push_back(':');
  const auto r = std::to_chars(buf, buf + sizeof(buf), v ? v : kEps, std::chars_format::fixed, 6);
  out.append(buf, r.ptr);
}

// ":number" -> ":%f" of it, appended to `out`; returns the pointer behind the number. The common spelling - up to
// three integer digits, a point, digits - is rounded to six places on the digits themselves: "%f" rounds the exact
// value of the nearest double, which lies within 1000 * 2^-53 = 1.2e-13 of the decimal text, so the two roundings
// can only differ when the discarded digits are that close to a tie - they start "5" + five zeros or "4" + five
// nines - and those (and zero, which becomes EPS, and every other spelling) take the parse-and-print path.
inline const char* nw_copy_length(const char* p, const char* end, std::string& out) {
  const char* q = nw_skip(p);
  if (*q == ':') {
    const char* b = nw_skip(q + 1);
    const char* s = b;
    while (*s >= '0' && *s <= '9') ++s;
    const int ni = (int)(s - b);
    if (ni >= 1 && ni <= 3 && (ni == 1 || *b != '0') && *s == '.') {
      const char* f = s + 1;
      const char* e = f;
      while (*e >= '0' && *e <= '9') ++e;
      const int nf = (int)(e - f);
      if (nf >= 1 && nw_delim(*e)) {
        bool nonzero = false;
        for (const char* c = b; c < e && !nonzero; ++c) nonzero = *c != '0' && *c != '.';
        bool decided = nonzero, up = false;
        if (decided && nf > 6) {
          const char r0 = f[6];
          if (r0 == '4' || r0 == '5') {
            const char near_tie = r0 == '4' ? '9' : '0';
            bool all = true;
            for (int k = 7; k < 12 && all; ++k) all = (k < nf ? f[k] : '0') == near_tie;
            decided = !all;
          }
          up = r0 >= '5';
        }
        if (decided) {
          char buf[16];
          int n = 0;
          buf[n++] = '0';   // room for a carry out of the integer part
          for (int k = 0; k < ni; ++k) buf[n++] = b[k];
          buf[n++] = '.';
          for (int k = 0; k < 6; ++k) buf[n++] = k < nf ? f[k] : '0';
          if (up) {
            int k = n - 1;
            for (;; --k) {
              if (buf[k] == '.') continue;
              if (buf[k] == '9') {
                buf[k] = '0';
              } else {
                ++buf[k];
                break;
              }
            }
          }
          out.push_back(':');
          const int from = buf[0] == '0' ? 1 : 0;
          out.append(buf + from, (std::size_t)(n - from));
          return e;
        }
      }
    }
  }
  double len;
  p = nw_length(p, end, &len);
  nw_put_length(out, len);
  return p;
}

}  // namespace

std::string ExportNewick(const std::string& newick) {
  const char* p = newick.c_str();
  const char* const end = p + newick.size();
  std::string out;
  out.reserve(newick.size());
  std::vector<int> kids;  // children finished so far, per open node
  bool want_item = true;
  for (;;) {
    p = nw_skip(p);
    if (want_item) {
      if (*p == '(') {
        out.push_back('(');
        kids.push_back(0);
        ++p;
        continue;
      }
      const char* b = p;
      while (!nw_delim(*p)) ++p;
      if (p == b) throw std::invalid_argument("empty tip label in Newick string");
      out.append(b, p);
      p = nw_copy_length(p, end, out);
      if (kids.empty()) return out;  // a single tip
      ++kids.back();
      want_item = false;
    } else if (*p == ',') {
      out.push_back(',');
      ++p;
      want_item = true;
    } else if (*p == ')') {
      ++p;
      out.push_back(')');
      p = nw_skip(p);
      const char* b = p;
      while (!nw_delim(*p)) ++p;
      out.append(b, p);
      const std::size_t before_length = out.size();
      p = nw_copy_length(p, end, out);   // (read first, as before: a malformed length is reported before the shape)
      const bool top = kids.size() == 1;
      if (kids.back() != (top ? 3 : 2)) throw std::invalid_argument("Newick tree is not an unrooted binary tree");
      kids.pop_back();
      if (top) {
        out.resize(before_length);       // a length on the trifurcation is dropped
        out.append(":0.0;");
        return out;
      }
      ++kids.back();
    } else {
      throw std::invalid_argument("unbalanced Newick string");
    }
  }
}

// flat arrays behind lh_family_desc_t
struct PhyloHMM::FlatFamily {
  std::vector<uint8_t> tip_codes;
  std::vector<int32_t> off[5], inds[5];
  std::vector<int32_t> vd_inds, dj_inds;
  std::vector<double> v_gene_prob, v_trans_prod;
  std::vector<const char*> label_ptrs;
  lh_family_desc_t desc;
};

PhyloHMM::PhyloHMM(const std::string& yaml_path, int cluster_ind, const std::string& hmm_param_dir, int seed)
    : HMM(yaml_path, cluster_ind, hmm_param_dir, seed) {
  InitializeXmsaStructs();
  BuildFlatFamily();
}

PhyloHMM::~PhyloHMM() {
  if (dev_group_) lh_group_destroy(dev_group_);   // (owns its members)
  else if (dev_family_) lh_family_destroy(dev_family_);
}

// src/PhyloHMM.cpp:45-89, 123-144
void PhyloHMM::InitializeXmsaStructs() {
  xmsa_labels_.clear();
  xmsa_labels_.push_back("naive");
  for (const auto& id : cluster_data_.at("unique_ids").seq) xmsa_labels_.push_back(id.as_string());
  xmsa_naive_ind_ = 0;

  XmsaIds ids;
  StoreGermlinePaddingXmsaIndices(vpadding_, ids);
  StoreGermlinePaddingXmsaIndices(vgerm_, ids);
  if (locus_ == "igh") {
    StoreJunctionXmsaIndices(vd_junction_, flexbounds_.at("v_r"), flexbounds_.at("d_l"), ids);
    StoreGermlinePaddingXmsaIndices(dgerm_, ids);
    StoreJunctionXmsaIndices(dj_junction_, flexbounds_.at("d_r"), flexbounds_.at("j_l"), ids);
  } else {
    StoreJunctionXmsaIndices(vd_junction_, flexbounds_.at("v_r"), flexbounds_.at("j_l"), ids);
  }
  StoreGermlinePaddingXmsaIndices(jgerm_, ids);
  StoreGermlinePaddingXmsaIndices(jpadding_, ids);

  // BuildXmsa: column = [naive base ; the MSA column of the site]
  const int n = msa_.rows();
  xmsa_.assign(n + 1, (int)ids.size(), -1);
  for (const auto& kv : ids) {
    const int naive_base = kv.first.first, site = kv.first.second, col = kv.second;
    if (site < 0 || site >= msa_.cols()) throw std::runtime_error("flexbounds / relpos reach outside the alignment");
    xmsa_(0, col) = naive_base;
    for (int i = 0; i < n; ++i) xmsa_(i + 1, col) = msa_(i, site);
  }
  xmsa_seqs_.assign(n + 1, std::string(xmsa_.cols(), ' '));
  for (int i = 0; i <= n; ++i)
    for (int j = 0; j < xmsa_.cols(); ++j) xmsa_seqs_[i][j] = alphabet_.at(xmsa_(i, j));
}

void PhyloHMM::BuildFlatFamily() {
  flat_.reset(new FlatFamily);
  FlatFamily& F = *flat_;
  const int T = xmsa_.rows(), S = xmsa_.cols();
  F.tip_codes.resize((std::size_t)T * S);
  for (int i = 0; i < T; ++i)
    for (int j = 0; j < S; ++j) F.tip_codes[(std::size_t)i * S + j] = (uint8_t)xmsa_(i, j);
  const GermlineRegion* regions[5] = {&vpadding_, &vgerm_, &dgerm_, &jgerm_, &jpadding_};
  for (int r = 0; r < 5; ++r) {
    const GermlineRegion& R = *regions[r];
    F.off[r].assign(1, 0);
    for (const auto& kv : R.ggene_ranges) {  // ranges are contiguous in map order
      for (int i = kv.second.first; i < kv.second.second; ++i) F.inds[r].push_back(R.xmsa_inds[i]);
      F.off[r].push_back((int32_t)F.inds[r].size());
    }
    if (F.inds[r].empty()) F.inds[r].push_back(0);
  }
  F.vd_inds.assign(vd_junction_.xmsa_inds.v.begin(), vd_junction_.xmsa_inds.v.end());
  F.dj_inds.assign(dj_junction_.xmsa_inds.v.begin(), dj_junction_.xmsa_inds.v.end());
  if (F.dj_inds.empty()) F.dj_inds.push_back(-1);
  for (const auto& kv : vgerm_.ggene_ranges) {  // src/HMM.cpp:296-314
    const lhb::GermlineGene& g = ggenes_.at(kv.first);
    F.v_gene_prob.push_back(g.gene_prob);
    // product of the in-region germline transitions in Eigen's reduction order (see hmm_model.cpp)
    const int s = vgerm_.germ_inds.at(kv.second.first), len = kv.second.second - kv.second.first - 1;
    const double prod = EigenOrderProduct(g.transition.data() + s, len, s);
    F.v_trans_prod.push_back(prod);
  }
  for (const std::string& l : xmsa_labels_) F.label_ptrs.push_back(l.c_str());

  lh_family_desc_t& d = F.desc;
  std::memset(&d, 0, sizeof(d));
  d.n_tips = T;
  d.n_sites = S;
  d.tip_codes = F.tip_codes.data();
  d.has_d = locus_ == "igh";
  d.n_v = (int)vgerm_.state_strs.size();
  d.n_d = (int)dgerm_.state_strs.size();
  d.n_j = (int)jgerm_.state_strs.size();
  d.vd_sites = vd_junction_.xmsa_inds.rows();
  d.vd_states = vd_junction_.xmsa_inds.cols();
  d.dj_sites = dj_junction_.xmsa_inds.rows();
  d.dj_states = dj_junction_.xmsa_inds.cols();
  d.vpad_off = F.off[0].data();  d.vpad_inds = F.inds[0].data();
  d.vgerm_off = F.off[1].data(); d.vgerm_inds = F.inds[1].data();
  d.dgerm_off = F.off[2].data(); d.dgerm_inds = F.inds[2].data();
  d.jgerm_off = F.off[3].data(); d.jgerm_inds = F.inds[3].data();
  d.jpad_off = F.off[4].data();  d.jpad_inds = F.inds[4].data();
  d.vd_inds = F.vd_inds.data();
  d.dj_inds = F.dj_inds.data();
  d.v_gene_prob = F.v_gene_prob.data();
  d.v_trans_prod = F.v_trans_prod.data();
  d.vpad_trans = vpadding_transition_.data();
  d.jpad_trans = jpadding_transition_.data();
  d.t_v_vd = vgerm_vd_junction_transition_.data();
  d.t_vd = vd_junction_transition_.data();
  d.t_vd_d = vd_junction_dgerm_transition_.data();
  d.t_d_dj = dgerm_dj_junction_transition_.data();
  d.t_dj = dj_junction_transition_.data();
  d.t_dj_j = dj_junction_jgerm_transition_.data();
}

const lh_family_desc& PhyloHMM::family_desc() const { return flat_->desc; }

void PhyloHMM::EnsureDeviceFamily() {
  if (dev_family_) return;
  if (devices_.size() > 1) {
    if (lh_group_create(&dev_group_, &flat_->desc, devices_.data(), (int)devices_.size())) throw std::runtime_error(lh_group_last_error());
    dev_family_ = lh_group_member(dev_group_, 0);
    if (host_threads_ > 0 && lh_group_set_option(dev_group_, "host_threads", std::max(1, host_threads_ / (int)devices_.size())))
      throw std::runtime_error(lh_group_last_error());
    if (lh_group_set_labels(dev_group_, flat_->label_ptrs.data())) throw std::runtime_error(lh_group_last_error());
    return;
  }
  if (lh_family_create(&dev_family_, &flat_->desc, devices_[0])) throw std::runtime_error(lh_last_error());
  if (host_threads_ > 0 && lh_set_option(dev_family_, "host_threads", host_threads_)) throw std::runtime_error(lh_last_error());
  // the xMSA row labels, for the device Newick parser (Partition::SetTipStates matches tips by label,
  // lib/libptpll/src/pll_partition.cpp:153-194)
  if (lh_family_set_labels(dev_family_, flat_->label_ptrs.data())) throw std::runtime_error(lh_last_error());
}

void PhyloHMM::set_devices(const std::vector<int>& devices) {
  if (devices.empty()) throw std::invalid_argument("empty device list");
  if (dev_family_) throw std::logic_error("set_devices() must come before the first scoring call");
  devices_ = devices;
}

void PhyloHMM::set_host_threads(int n) {
  if (n < 0) throw std::invalid_argument("host thread count must not be negative");
  host_threads_ = n;
  if (dev_group_ && n > 0) {
    if (lh_group_set_option(dev_group_, "host_threads", std::max(1, n / (int)devices_.size()))) throw std::runtime_error(lh_group_last_error());
  } else if (dev_family_ && n > 0 && lh_set_option(dev_family_, "host_threads", n)) {
    throw std::runtime_error(lh_last_error());
  }
}

// src/PhyloHMM.cpp:350-361
void PhyloHMM::InitializePhyloParameters(const std::string& newick_path, const std::vector<double>& er,
                                         const std::vector<double>& pi, double alpha, int num_rates) {
  if (er.size() != 6 || pi.size() != 4) throw std::invalid_argument("er must have 6 and pi 4 entries");
  tree_ = read_file(newick_path);
  forward_ready_ = false;
  er_ = er;
  pi_ = pi;
  alpha_ = alpha;
  sr_.assign(num_rates, 0.0);
  if (lhb::gamma_mean_rates(alpha_, num_rates, sr_.data())) throw std::invalid_argument(lhh_last_error());
}

// One device call = Partition construction + pruning + emission fill (InitializePhyloEmission,
// src/PhyloHMM.cpp:366-383) + the forward pass (RunForwardAlgorithm, src/HMM.cpp:254-287).
void PhyloHMM::ScoreCurrentSample() {
  EnsureDeviceFamily();
  const int T = xmsa_.rows(), R = (int)sr_.size();
  std::vector<int32_t> children((std::size_t)2 * (T - 2)), root_child(1);
  std::vector<double> blen((std::size_t)2 * (T - 2)), root_blen(1);
  const char* nw = tree_.c_str();
  if (lhh_newick_to_ops(T, flat_->label_ptrs.data(), 1, &nw, 1, children.data(), blen.data(), root_child.data(), root_blen.data()))
    throw std::invalid_argument(lhh_last_error());
  lh_tree_batch_t tb{1, R, children.data(), blen.data(), root_child.data(), root_blen.data(), er_.data(), pi_.data(), sr_.data()};
  const lh_family_desc_t& d = flat_->desc;
  xmsa_emission_.assign(d.n_sites, 0.0);
  vgerm_.forward.assign(d.n_v, 0.0);
  dgerm_.forward.assign(d.n_d, 0.0);
  jgerm_.forward.assign(d.n_j, 0.0);
  vd_junction_.forward.assign(d.vd_sites, d.vd_states, 0.0);
  dj_junction_.forward.assign(d.dj_sites, d.dj_states, 0.0);
  vd_junction_.scaler_counts.assign(d.vd_sites, 0);
  dj_junction_.scaler_counts.assign(d.dj_sites, 0);
  int32_t gs[3] = {0, 0, 0};
  std::vector<double> dgerm_dummy(1), dj_dummy(1);
  std::vector<int32_t> djs_dummy(1);
  lh_outputs_t o;
  std::memset(&o, 0, sizeof(o));
  o.xmsa_emission = xmsa_emission_.data();
  o.vgerm_forward = vgerm_.forward.data();
  o.vd_forward = vd_junction_.forward.data();
  o.dgerm_forward = d.has_d ? dgerm_.forward.data() : dgerm_dummy.data();
  o.dj_forward = d.has_d ? dj_junction_.forward.data() : dj_dummy.data();
  o.jgerm_forward = jgerm_.forward.data();
  o.germ_scalers = gs;
  o.vd_scalers = vd_junction_.scaler_counts.data();
  o.dj_scalers = d.has_d ? dj_junction_.scaler_counts.data() : djs_dummy.data();
  double ll = 0.0;
  if (lh_score_batch(dev_family_, &tb, &ll, &o)) throw std::runtime_error(lh_last_error());
  vgerm_.scaler_count = gs[0];
  dgerm_.scaler_count = gs[1];
  jgerm_.scaler_count = gs[2];
  loglik_cached_ = ll;
  cache_forward_ = false;
  forward_ready_ = true;
}

// src/PhyloHMM.cpp:366-383. The device call fills the emissions and runs the forward pass in one go; the flag a
// caller can observe still says, as in the reference, that the forward pass is due (cache_forward() == true until
// the first LogLikelihood / SampleNaiveSequence, test/test.cpp:966), and RunForwardAlgorithm then finds it done.
void PhyloHMM::InitializePhyloEmission() {
  if (tree_.empty()) throw std::runtime_error("InitializePhyloEmission() called before a tree was set");
  ScoreCurrentSample();
  cache_forward_ = true;
}

void PhyloHMM::RunForwardAlgorithm() {
  if (!forward_ready_) ScoreCurrentSample();
}

// src/HMM.cpp:345-354
double PhyloHMM::LogLikelihood() {
  if (cache_forward_) {
    RunForwardAlgorithm();
    cache_forward_ = false;
  }
  return loglik_cached_;
}

std::vector<double> PhyloHMM::ScoreSamples(const std::vector<std::string>& newicks, const std::vector<double>& er,
                                           const std::vector<double>& pi, const std::vector<double>& alpha, int num_rates) {
  const int B = (int)newicks.size();
  if ((int)er.size() != 6 * B || (int)pi.size() != 4 * B || (int)alpha.size() != B)
    throw std::invalid_argument("ScoreSamples: er / pi / alpha do not match the number of trees");
  std::vector<const char*> nw(B);
  for (int b = 0; b < B; ++b) nw[b] = newicks[b].c_str();
  std::vector<double> ll(B);
  ScoreSamples(B, nw.data(), er.data(), pi.data(), alpha.data(), num_rates, ll.data());
  return ll;
}

// The loop body src/PhyloHMM.cpp:419-430 for B rows of the RevBayes trace in one device call: the Newick
// strings go to the GPU as text and are parsed there (k_newick), the discrete-Gamma rates are computed on
// the handle's host threads (bit-identical to pll_compute_gamma_cats), planning, pruning and the forward
// pass follow chunk by chunk while later strings are still being copied.
void PhyloHMM::ScoreSamples(int B, const char* const* newicks, const double* er, const double* pi, const double* alpha,
                            int num_rates, double* loglik) {
  ScoreSamples(B, newicks, nullptr, er, pi, alpha, num_rates, loglik);
}

void PhyloHMM::ScoreSamples(int B, const char* const* newicks, const int64_t* lengths, const double* er, const double* pi,
                            const double* alpha, int num_rates, double* loglik) {
  EnsureDeviceFamily();
  if (B <= 0) return;
  lh_newick_batch_t nb{};
  nb.n_trees = B;
  nb.n_rates = num_rates;
  nb.newick = newicks;
  nb.length = lengths;
  nb.er = er;
  nb.pi = pi;
  nb.alpha = alpha;
  const int rc = dev_group_ ? lh_group_score_newick(dev_group_, &nb, loglik, nullptr) : lh_score_newick(dev_family_, &nb, loglik, nullptr);
  if (rc) {
    const std::string msg = dev_group_ ? lh_group_last_error() : lh_last_error();
    // what the reference's parser / SetTipStates would refuse is the caller's mistake, the rest the device's
    if (msg.rfind("tree ", 0) == 0 || msg.find("alpha") != std::string::npos) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
  }
}

// src/PhyloHMM.cpp:244-282
void PhyloHMM::WriteOutputHeaders(std::ostream& out) const {
  out << "Iteration\tRBLogLikelihood\tPrior\talpha\t";
  for (std::size_t i = 1; i <= er_.size(); ++i) out << "er[" << i << "]\t";
  for (std::size_t i = 1; i <= pi_.size(); ++i) out << "pi[" << i << "]\t";
  out << "tree\t";
  for (std::size_t i = 1; i <= sr_.size(); ++i) out << "sr[" << i << "]\t";
  out << "LHLogLikelihood\tLogWeight\tNaiveSequence\tVGene\tV5pDel\tV3pDel\tVFwkInsertion\t";
  if (locus_ == "igh") out << "VDInsertion\tDGene\tD5pDel\tD3pDel\tDJInsertion\t";
  else out << "VJInsertion\t";
  out << "JGene\tJ5pDel\tJ3pDel\tJFwkInsertion\n";
}

// src/PhyloHMM.cpp:288-327. One output line from the row's values and the annotation `h` sampled last
// (a free function so that RunPipeline's sampling clones format their rows with the very same code).
namespace {
struct LineFields {
  int iteration;
  double rb, prior, alpha;
  const double *er, *pi, *sr;
  int n_er, n_pi, n_sr;
  const std::string* tree_text;   // already exported (pll_utree_export_newick)
  double lh, logweight;
  const std::string* naive;
};
void write_output_line(std::ostream& out, const LineFields& f, const HMM& h) {
  out << f.iteration << "\t" << f.rb << "\t" << f.prior << "\t" << f.alpha << "\t";
  for (int i = 0; i < f.n_er; ++i) out << f.er[i] << "\t";
  for (int i = 0; i < f.n_pi; ++i) out << f.pi[i] << "\t";
  out << *f.tree_text << "\t";
  for (int i = 0; i < f.n_sr; ++i) out << f.sr[i] << "\t";
  out << f.lh << "\t" << f.logweight << "\t" << *f.naive << "\t";
  out << h.vgerm_state_str_samp() << "\t" << h.vgerm_left_del_samp() << "\t" << h.vgerm_right_del_samp() << "\t"
      << h.vgerm_left_insertion_samp() << "\t";
  if (h.locus() == "igh") {
    out << h.vd_junction_insertion_samp() << "\t" << h.dgerm_state_str_samp() << "\t" << h.dgerm_left_del_samp() << "\t"
        << h.dgerm_right_del_samp() << "\t" << h.dj_junction_insertion_samp() << "\t";
  } else {
    out << h.vd_junction_insertion_samp() << "\t";
  }
  out << h.jgerm_state_str_samp() << "\t" << h.jgerm_left_del_samp() << "\t" << h.jgerm_right_del_samp() << "\t"
      << h.jgerm_right_insertion_samp() << "\n";
}
}  // namespace

void PhyloHMM::WriteOutputLine(std::ostream& out) const {
  const std::string exported = tree_export_cache_.empty() ? ExportNewick(tree_) : tree_export_cache_;
  const LineFields f{iteration_, rb_loglikelihood_, prior_, alpha_, er_.data(), pi_.data(), sr_.data(), (int)er_.size(),
                     (int)pi_.size(), (int)sr_.size(), &exported, lh_loglikelihood_, logweight_, &naive_sequence_};
  write_output_line(out, f, *this);
}

// src/PhyloHMM.cpp:393-446, batched and pipelined. Three stages run side by side over a ring of three batches:
//   reader  (a thread)   reads up to batch_size rows of the RevBayes trace, converts their numbers, computes the
//                        discrete-Gamma rates (bit-identical to pll_compute_gamma_cats) and the exported tree text
//   scorer  (a thread)   one device call per batch, Newick text in (parsed on the GPU), forward tables out
//   sampler (the caller) finishes every row in file order: SampleNaiveSequence + the output line. The RNG
//                        advances exactly as in the reference: one std::mt19937 walked row by row
//                        (src/HMM.cpp:325-329, 1252-1257).
// So the TSV read / ingest of batch b+1, the device scoring of b and the host sampling of b-1 overlap.
namespace {
struct PipeRow {
  int iteration;
  double rb, prior, alpha, er[6], pi[4];
  std::string tree, exported;
};
struct PipeBatch {
  std::vector<PipeRow> rows;
  std::vector<const char*> nw;
  std::vector<int64_t> len;
  std::vector<double> er, pi, alpha, rates, ll, xe, fv, fvd, fd, fdj, fj;
  std::vector<int32_t> gs, vds, djs;
  std::string err;          // a stage's failure, reported by the sampler in file order
  bool err_is_input = false;
  int state = 0;            // 0 free, 1 read, 2 scored
  // the draws of the rows' annotations made on the device (lh_sample_paths), or where the run's generator stood
  // before the batch for the host sampler
  std::vector<int32_t> idx;
  std::vector<uint8_t> sflag;
  std::mt19937 rng_start;
  bool device_sampled = false;
};

// A copy of what SampleNaiveSequence reads and writes - state space, transitions, forward tables, the sampled
// annotation, the RNG - so that RunPipeline can sample the rows of a batch on several threads. The reference draws
// every row from ONE std::mt19937 in file order (src/HMM.cpp:325-329, 1252-1257); a row consumes a fixed number of
// its outputs (one std::discrete_distribution draw per germline region and junction site, each over a weight vector
// whose length the family fixes), so the generator's state at the start of row i is the batch's start state advanced
// by i times that number (std::mt19937::discard), and the rows come out exactly as the serial walk gives them. The
// block ends are checked against the next block's start; a mismatch sends the batch through the serial path.
class SamplerClone : public HMM {
 public:
  explicit SamplerClone(const HMM& h) : HMM(h) { cache_forward_ = false; }
  double LogLikelihood() override { return 0.0; }
  std::mt19937& rng() { return rng_; }
  void Load(const PipeBatch& b, int i, const lh_family_desc_t& d) {
    const std::size_t vd_n = (std::size_t)d.vd_sites * d.vd_states, dj_n = (std::size_t)d.dj_sites * d.dj_states;
    vgerm_.forward.assign(b.fv.begin() + (std::size_t)i * d.n_v, b.fv.begin() + (std::size_t)(i + 1) * d.n_v);
    jgerm_.forward.assign(b.fj.begin() + (std::size_t)i * d.n_j, b.fj.begin() + (std::size_t)(i + 1) * d.n_j);
    vd_junction_.forward.r = d.vd_sites;
    vd_junction_.forward.c = d.vd_states;
    vd_junction_.forward.v.assign(b.fvd.begin() + i * vd_n, b.fvd.begin() + (i + 1) * vd_n);
    if (d.has_d) {
      dgerm_.forward.assign(b.fd.begin() + (std::size_t)i * d.n_d, b.fd.begin() + (std::size_t)(i + 1) * d.n_d);
      dj_junction_.forward.r = d.dj_sites;
      dj_junction_.forward.c = d.dj_states;
      dj_junction_.forward.v.assign(b.fdj.begin() + i * dj_n, b.fdj.begin() + (i + 1) * dj_n);
    }
    cache_forward_ = false;
  }

 protected:
  void RunForwardAlgorithm() override {}
};

// outputs of the generator one row's sampling consumes: libstdc++'s discrete_distribution draws one
// generate_canonical<double, 53> (two 32-bit outputs) unless it has fewer than two weights (then none)
unsigned long long rng_outputs_per_row(const lh_family_desc_t& d) {
  auto draws = [](int n_weights, int calls) { return n_weights >= 2 ? 2ull * calls : 0ull; };
  unsigned long long n = draws(d.n_j, 1) + draws(d.vd_states, d.vd_sites) + draws(d.n_v, 1);
  if (d.has_d) n += draws(d.dj_states, d.dj_sites) + draws(d.n_d, 1);
  return n;
}
}  // namespace

void PhyloHMM::RunPipeline(const std::string& input_path, const std::string& output_path, int num_rates, int batch_size) {
  if (batch_size < 1) throw std::invalid_argument("batch_size must be positive");
  std::ifstream in(input_path);
  if (!in) throw std::runtime_error("cannot open " + input_path);
  std::string header_line;
  if (!std::getline(in, header_line)) throw std::runtime_error(input_path + " is empty");
  const std::vector<std::string> header = split_tsv(header_line);
  static const char* kCols[15] = {"Iteration", "Likelihood", "Prior", "alpha", "er[1]", "er[2]", "er[3]", "er[4]",
                                  "er[5]", "er[6]", "pi[1]", "pi[2]", "pi[3]", "pi[4]", "tree"};
  int col[15];
  for (int k = 0; k < 15; ++k) {
    col[k] = -1;
    for (std::size_t j = 0; j < header.size(); ++j)
      if (header[j] == kCols[k]) col[k] = (int)j;
    if (col[k] < 0) throw std::runtime_error(std::string("missing column \"") + kCols[k] + "\" in " + input_path);
  }
  std::ofstream outfile(output_path);
  if (!outfile) throw std::runtime_error("cannot open " + output_path);
  EnsureDeviceFamily();
  const lh_family_desc_t& d = flat_->desc;
  const std::size_t vd_n = (std::size_t)d.vd_sites * d.vd_states, dj_n = (std::size_t)d.dj_sites * d.dj_states;
  sr_.assign(num_rates, 0.0);
  // The draws of SampleNaiveSequence are made on the device, from the forward tables k_hmm leaves there
  // (lh_sample_paths), when the run is on one device: the host then only turns the drawn states into the annotation
  // strings, and the tables (53 KB per row at the benchmark shapes) stay where they are. The one generator of the
  // run is advanced by the scorer thread in file order - a fixed number of raw outputs per row - and handed back at
  // the end. LH_HOST_SAMPLER=1 (or several devices) keeps the host sampler.
  const unsigned long long words_per_row = rng_outputs_per_row(d);
  const int n_idx = 2 + d.vd_sites + (d.has_d ? 1 + d.dj_sites : 0);
  const bool dev_sampler = !dev_group_ && device_sampler_ && words_per_row > 0 && !std::getenv("LH_HOST_SAMPLER");
  std::mt19937 run_gen = rng_;
  if (dev_sampler && lh_set_option(dev_family_, "keep_tables", 1)) throw std::runtime_error(lh_last_error());
  struct KeepTablesOff {
    lh_family* f;
    ~KeepTablesOff() { if (f) lh_set_option(f, "keep_tables", 0); }
  } keep_tables_off{dev_sampler ? dev_family_ : nullptr};

  constexpr int kRing = 3;
  PipeBatch ring[kRing];
  std::mutex m;
  std::condition_variable cv;
  int n_read = 0;           // batches the reader has handed over; -1 - n: the input ended after n batches
  bool input_done = false, abort = false;
  auto wait_for = [&](PipeBatch& b, int state) {  // false: the pipeline is being torn down
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [&] { return abort || b.state == state; });
    return !abort;
  };
  auto hand_over = [&](PipeBatch& b, int state) {
    {
      std::lock_guard<std::mutex> lk(m);
      b.state = state;
    }
    cv.notify_all();
  };

  // ---- reader -----------------------------------------------------------------------------------------------
  double ms_reader = 0.0, ms_scorer = 0.0, ms_sampler = 0.0;   // busy time of the three stages (LH_TRACE)
  const int hw = host_threads_ > 0 ? host_threads_ : (int)std::max(1u, std::thread::hardware_concurrency());
  auto env_int = [](const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v && *v ? std::max(1, std::atoi(v)) : dflt;
  };
  // (with the draws on the device the rows' TSV fields and tree export are most of the host's work: the reader gets
  // three quarters of the threads, the writer of the output lines a quarter; with the host sampler half and half)
  const int reader_threads = env_int("LH_READER_THREADS", std::max(1, dev_sampler ? hw * 3 / 4 : hw / 2));
  const int sampler_threads = env_int("LH_SAMPLER_THREADS", std::max(1, std::min(dev_sampler ? hw / 4 : hw / 2, 16)));
  // (persistent workers for the two host stages: a thread created and joined per batch of 256 rows costs as much as
  // a few rows)
  lhb::WorkerPool reader_pool(std::min(reader_threads, 16) - 1), writer_pool(std::max(1, sampler_threads) - 1);
  std::thread reader([&] {
    std::string line;
    std::vector<std::string> lines;
    bool eof = false;
    // the file comes in blocks of 4 MB and is cut at the newlines with memchr (std::getline on the stream moved a
    // 9 KB row at about 1 GB/s on one thread - as long as all the other work of the row on eight)
    std::string pending;
    std::size_t ppos = 0;
    bool file_end = false;
    constexpr std::size_t kReadBlock = (std::size_t)4 << 20;
    auto next_line = [&](std::string& out) -> bool {
      for (;;) {
        const char* base = pending.data();
        const void* nl = ppos < pending.size() ? std::memchr(base + ppos, '\n', pending.size() - ppos) : nullptr;
        if (nl) {
          const std::size_t e = (std::size_t)(static_cast<const char*>(nl) - base);
          out.assign(base + ppos, e - ppos);
          ppos = e + 1;
          return true;
        }
        if (file_end) {
          if (ppos >= pending.size()) return false;
          out.assign(base + ppos, pending.size() - ppos);   // a last line without a newline
          ppos = pending.size();
          return true;
        }
        pending.erase(0, ppos);
        ppos = 0;
        const std::size_t old = pending.size();
        pending.resize(old + kReadBlock);
        in.read(&pending[old], (std::streamsize)kReadBlock);
        const std::size_t got = (std::size_t)in.gcount();
        pending.resize(old + got);
        if (got < kReadBlock) file_end = true;
      }
    };
    for (int bi = 0; !eof; ++bi) {
      PipeBatch& b = ring[bi % kRing];
      if (!wait_for(b, 0)) return;
      b.rows.clear();
      b.err.clear();
      try {
        const auto t_r0 = std::chrono::steady_clock::now();
        // the lines are read one after the other; their fields (15 numbers, a tree of several KB to re-export)
        // are independent and are worked on by a few threads
        lines.clear();
        // (the first batches are short - 128 rows, then twice as many each time up to batch_size - so that the
        // scorer and the sampler start after 128 rows have been read, not after batch_size: on a trace of a few
        // thousand rows the fill of a three-stage pipeline is otherwise a third of the run)
        const int want_rows = std::min<long long>(batch_size, 128LL << std::min(bi, 20));
        while ((int)lines.size() < want_rows) {
          if (!next_line(line)) { eof = true; break; }
          if (line.empty()) continue;
          lines.emplace_back(std::move(line));
        }
        const int n_lines = (int)lines.size();
        b.rows.assign(n_lines, PipeRow{});
        std::vector<std::string> row_err(n_lines);
        auto parse_rows = [&](int w, int nw) {
          for (int i = w; i < n_lines; i += nw) {
            try {
              std::vector<std::string> f = split_tsv(lines[i]);
              if ((int)f.size() < (int)header.size()) throw std::runtime_error("too few columns in a row of " + input_path);
              PipeRow& r = b.rows[i];
              r.iteration = (int)to_double(f[col[0]], "Iteration");
              r.rb = to_double(f[col[1]], "Likelihood");
              r.prior = to_double(f[col[2]], "Prior");
              r.alpha = to_double(f[col[3]], "alpha");
              for (int k = 0; k < 6; ++k) r.er[k] = to_double(f[col[4 + k]], "er");
              for (int k = 0; k < 4; ++k) r.pi[k] = to_double(f[col[10 + k]], "pi");
              r.tree = std::move(f[col[14]]);
              // the exported tree text of the output line (pll_utree_export_newick of the parsed tree; text work
              // only). A string it cannot read is left to the device call, whose parser words the error.
              try {
                r.exported = ExportNewick(r.tree);
              } catch (const std::exception&) {
                r.exported.clear();
              }
            } catch (const std::exception& e) {
              row_err[i] = e.what();
            }
          }
        };
        const int n_workers = std::max(1, std::min({reader_threads, n_lines / 16, 16}));
        if (n_workers > 1) reader_pool.run(n_workers, [&](int w, int) { parse_rows(w, n_workers); });
        else parse_rows(0, 1);
        for (int i = 0; i < n_lines; ++i)
          if (!row_err[i].empty()) throw std::runtime_error(row_err[i]);   // the first bad row, in file order
        ms_reader += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_r0).count();
        const std::size_t B = b.rows.size();
        b.nw.resize(B);
        b.len.resize(B);
        b.er.resize(6 * B);
        b.pi.resize(4 * B);
        b.alpha.resize(B);
        b.rates.resize(B * (std::size_t)num_rates);
        for (std::size_t i = 0; i < B; ++i) {
          b.nw[i] = b.rows[i].tree.c_str();
          b.len[i] = (int64_t)b.rows[i].tree.size();
          std::memcpy(&b.er[6 * i], b.rows[i].er, sizeof(double) * 6);
          std::memcpy(&b.pi[4 * i], b.rows[i].pi, sizeof(double) * 4);
          b.alpha[i] = b.rows[i].alpha;
        }
        if (B && lhh_gamma_cats_batch((int)B, b.alpha.data(), num_rates, b.rates.data())) {
          b.err = lhh_last_error();
          b.err_is_input = true;
        }
      } catch (const std::exception& e) {
        b.err = e.what();
        eof = true;
      }
      {
        std::lock_guard<std::mutex> lk(m);
        b.state = 1;
        ++n_read;
        if (eof) input_done = true;
      }
      cv.notify_all();
    }
  });

  // ---- scorer -----------------------------------------------------------------------------------------------
  std::thread scorer([&] {
    std::vector<uint32_t> words;   // the generator's raw outputs for one batch, in file order
    for (int bi = 0;; ++bi) {
      {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return abort || n_read > bi || input_done; });
        if (abort || n_read <= bi) return;   // torn down, or the input ended before this batch
      }
      PipeBatch& b = ring[bi % kRing];
      const std::size_t B = b.rows.size();
      b.device_sampled = false;
      if (B && b.err.empty()) {
        b.ll.resize(B);
        b.xe.resize(B * (std::size_t)d.n_sites);
        b.fv.resize(B * (std::size_t)d.n_v);
        b.fvd.resize(B * vd_n);
        b.fd.resize(B * (std::size_t)std::max(d.n_d, 1));
        b.fdj.resize(B * std::max<std::size_t>(dj_n, 1));
        b.fj.resize(B * (std::size_t)d.n_j);
        b.gs.resize(B * 3);
        b.vds.resize(B * (std::size_t)d.vd_sites);
        b.djs.resize(B * (std::size_t)std::max(d.dj_sites, 1));
        lh_newick_batch_t nb{};
        nb.n_trees = (int)B;
        nb.n_rates = num_rates;
        nb.newick = b.nw.data();
        nb.length = b.len.data();
        nb.er = b.er.data();
        nb.pi = b.pi.data();
        nb.rates = b.rates.data();
        lh_outputs_t o{b.xe.data(), b.fv.data(), b.fvd.data(), b.fd.data(), b.fdj.data(), b.fj.data(),
                       b.gs.data(), b.vds.data(), b.djs.data(), nullptr};
        const auto t_s0 = std::chrono::steady_clock::now();
        int rc;
        b.device_sampled = false;
        if (dev_sampler) {
          b.rng_start = run_gen;
          rc = lh_score_newick(dev_family_, &nb, b.ll.data(), nullptr);
          if (!rc) {
            lh_timings_t tm;
            bool ok = lh_get_timings(dev_family_, &tm) == 0 && tm.newick_host_trees == 0;   // (else the tables are a sub-batch's)
            if (ok) {
              words.resize(B * words_per_row);
              for (uint32_t& w : words) w = (uint32_t)run_gen();
              b.idx.resize(B * (std::size_t)n_idx);
              b.sflag.resize(B);
              ok = lh_sample_paths(dev_family_, (int32_t)B, words.data(), (int32_t)words_per_row, b.idx.data(), n_idx,
                                   b.sflag.data()) == 0;
              for (std::size_t i = 0; ok && i < B; ++i) ok = b.sflag[i] == 0;
            } else {
              run_gen.discard(words_per_row * B);
            }
            b.device_sampled = ok;
            // (a batch the device leaves to the host - a non-finite forward entry somewhere - is scored once more
            // with the tables fetched; the host sampler starts from rng_start and uses the same outputs)
            if (!ok) rc = lh_score_newick(dev_family_, &nb, b.ll.data(), &o);
          }
        } else {
          rc = dev_group_ ? lh_group_score_newick(dev_group_, &nb, b.ll.data(), &o)
                          : lh_score_newick(dev_family_, &nb, b.ll.data(), &o);
        }
        ms_scorer += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_s0).count();
        if (rc) {
          b.err = dev_group_ ? lh_group_last_error() : lh_last_error();
          b.err_is_input = b.err.rfind("tree ", 0) == 0;   // what the reference's parser / SetTipStates would refuse
        }
      }
      hand_over(b, 2);
    }
  });

  // ---- sampler (this thread), in file order -------------------------------------------------------------------
  std::string failure;
  bool failure_is_input = false, first = true;
  const unsigned long long per_row = rng_outputs_per_row(d);
  std::vector<std::unique_ptr<SamplerClone>> clones;
  int last_clone = -1, last_batch = -1;   // where the last row's sampled state lives when a clone sampled it
  bool last_on_device = false;            // ... and its forward tables never came to the host
  try {
    for (int bi = 0;; ++bi) {
      {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return ring[bi % kRing].state == 2 || (input_done && n_read <= bi); });
        if (ring[bi % kRing].state != 2) break;   // the input ended before this batch
      }
      PipeBatch& b = ring[bi % kRing];
      if (!b.err.empty()) {
        failure = b.err;
        failure_is_input = b.err_is_input;
        break;
      }
      const int B = (int)b.rows.size();
      const auto t_w0 = std::chrono::steady_clock::now();
      // ---- the rows of the batch on several threads (see SamplerClone); the lines are written in file order ----
      bool sampled = false;
      const int n_blocks = std::min(sampler_threads, B / 32);
      if (b.device_sampled) {
        // the states are drawn (k_sample): what is left per row is the annotation's strings and the output line
        const int nb = std::max(1, n_blocks);
        while ((int)clones.size() < nb) clones.emplace_back(new SamplerClone(*this));
        std::vector<std::string> text(nb);
        const int per_block = (B + nb - 1) / nb;
        auto write_block = [&](int t) {
          const int lo = std::min(B, t * per_block), hi = std::min(B, lo + per_block);
          SamplerClone& c = *clones[t];
          std::ostringstream os;
          for (int i = lo; i < hi; ++i) {
            const PipeRow& r = b.rows[i];
            const std::string naive = c.ApplySampledPath(&b.idx[(std::size_t)i * n_idx]);
            const std::string exported = r.exported.empty() ? ExportNewick(r.tree) : r.exported;
            const LineFields f{r.iteration, r.rb, r.prior, r.alpha, r.er, r.pi, &b.rates[(std::size_t)i * num_rates], 6, 4, num_rates,
                               &exported, b.ll[i], b.ll[i] - r.rb, &naive};
            write_output_line(os, f, c);
          }
          text[t] = os.str();
        };
        if (nb > 1) writer_pool.run(nb, [&](int t, int) { write_block(t); });
        else write_block(0);
        if (first) {
          er_.assign(b.rows[0].er, b.rows[0].er + 6);
          pi_.assign(b.rows[0].pi, b.rows[0].pi + 4);
          WriteOutputHeaders(outfile);
          first = false;
        }
        for (const std::string& block : text) outfile << block;
        // (which clone holds the last row's annotation: the last block that is not empty)
        last_clone = std::min(nb - 1, (B - 1) / per_block);
        last_batch = bi % kRing;
        last_on_device = true;
        sampled = true;
      } else if (dev_sampler) {
        rng_ = b.rng_start;   // the host sampler below draws this batch's rows from where the generator stood
      }
      if (!sampled && n_blocks >= 2 && per_row > 0) {
        while ((int)clones.size() < n_blocks) clones.emplace_back(new SamplerClone(*this));
        std::vector<std::string> text(n_blocks);
        std::vector<std::mt19937> start(n_blocks + 1, rng_), finish(n_blocks, rng_);
        const int per_block = (B + n_blocks - 1) / n_blocks;
        auto sample_block = [&](int t) {
          const int lo = std::min(B, t * per_block), hi = std::min(B, lo + per_block);
          start[t].discard(per_row * (unsigned long long)lo);
          SamplerClone& c = *clones[t];
          c.rng() = start[t];
          std::ostringstream os;
          for (int i = lo; i < hi; ++i) {
            const PipeRow& r = b.rows[i];
            c.Load(b, i, d);
            const std::string naive = c.SampleNaiveSequence();
            const std::string exported = r.exported.empty() ? ExportNewick(r.tree) : r.exported;
            const LineFields f{r.iteration, r.rb, r.prior, r.alpha, r.er, r.pi, &b.rates[(std::size_t)i * num_rates], 6, 4, num_rates,
                               &exported, b.ll[i], b.ll[i] - r.rb, &naive};
            write_output_line(os, f, c);
          }
          finish[t] = c.rng();
          text[t] = os.str();
        };
        writer_pool.run(n_blocks, [&](int t, int) { sample_block(t); });
        start[n_blocks].discard(per_row * (unsigned long long)B);
        sampled = true;
        for (int t = 0; t < n_blocks; ++t) sampled = sampled && finish[t] == start[t + 1];
        if (sampled) {
          if (first) {
            er_.assign(b.rows[0].er, b.rows[0].er + 6);   // (the header names one column per rate and frequency)
            pi_.assign(b.rows[0].pi, b.rows[0].pi + 4);
            WriteOutputHeaders(outfile);
            first = false;
          }
          for (const std::string& block : text) outfile << block;
          rng_ = finish[n_blocks - 1];
          last_clone = n_blocks - 1;
          last_batch = bi % kRing;
          last_on_device = false;
        }
      }
      if (!sampled) {
        if (B) {
          last_clone = -1;
          last_on_device = false;
        }
        for (int i = 0; i < B; ++i) {
          const PipeRow& r = b.rows[i];
          iteration_ = r.iteration;
          rb_loglikelihood_ = r.rb;
          prior_ = r.prior;
          alpha_ = r.alpha;
          er_.assign(r.er, r.er + 6);
          pi_.assign(r.pi, r.pi + 4);
          tree_ = r.tree;
          tree_export_cache_ = r.exported;
          sr_.assign(b.rates.begin() + (std::size_t)i * num_rates, b.rates.begin() + (std::size_t)(i + 1) * num_rates);
          xmsa_emission_.assign(b.xe.begin() + (std::size_t)i * d.n_sites, b.xe.begin() + (std::size_t)(i + 1) * d.n_sites);
          vgerm_.forward.assign(b.fv.begin() + (std::size_t)i * d.n_v, b.fv.begin() + (std::size_t)(i + 1) * d.n_v);
          jgerm_.forward.assign(b.fj.begin() + (std::size_t)i * d.n_j, b.fj.begin() + (std::size_t)(i + 1) * d.n_j);
          vd_junction_.forward.assign(d.vd_sites, d.vd_states, 0.0);
          std::memcpy(vd_junction_.forward.data(), &b.fvd[i * vd_n], vd_n * sizeof(double));
          vd_junction_.scaler_counts.assign(b.vds.begin() + (std::size_t)i * d.vd_sites, b.vds.begin() + (std::size_t)(i + 1) * d.vd_sites);
          if (d.has_d) {
            dgerm_.forward.assign(b.fd.begin() + (std::size_t)i * d.n_d, b.fd.begin() + (std::size_t)(i + 1) * d.n_d);
            dj_junction_.forward.assign(d.dj_sites, d.dj_states, 0.0);
            std::memcpy(dj_junction_.forward.data(), &b.fdj[i * dj_n], dj_n * sizeof(double));
            dj_junction_.scaler_counts.assign(b.djs.begin() + (std::size_t)i * d.dj_sites, b.djs.begin() + (std::size_t)(i + 1) * d.dj_sites);
          }
          vgerm_.scaler_count = b.gs[3 * i];
          dgerm_.scaler_count = b.gs[3 * i + 1];
          jgerm_.scaler_count = b.gs[3 * i + 2];
          loglik_cached_ = b.ll[i];
          cache_forward_ = false;
          lh_loglikelihood_ = b.ll[i];
          logweight_ = lh_loglikelihood_ - rb_loglikelihood_;
          naive_sequence_ = SampleNaiveSequence();
          if (first) {
            WriteOutputHeaders(outfile);
            first = false;
          }
          WriteOutputLine(outfile);
        }
      }
      ms_sampler += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_w0).count();
      hand_over(b, 0);
    }
  } catch (...) {
    {
      std::lock_guard<std::mutex> lk(m);
      abort = true;
    }
    cv.notify_all();
    reader.join();
    scorer.join();
    tree_export_cache_.clear();
    throw;
  }
  {
    std::lock_guard<std::mutex> lk(m);
    abort = true;
  }
  cv.notify_all();
  reader.join();
  scorer.join();
  if (last_clone >= 0) {
    // the object ends as the reference's does, holding the last row: its forward tables and sampled annotation come
    // from the clone that sampled it (with the generator where the walk ended), the rest from the batch
    const PipeBatch& b = ring[last_batch];
    const int i = (int)b.rows.size() - 1;
    const PipeRow& r = b.rows[i];
    static_cast<HMM&>(*this) = *clones[last_clone];
    iteration_ = r.iteration;
    rb_loglikelihood_ = r.rb;
    prior_ = r.prior;
    alpha_ = r.alpha;
    er_.assign(r.er, r.er + 6);
    pi_.assign(r.pi, r.pi + 4);
    sr_.assign(b.rates.begin() + (std::size_t)i * num_rates, b.rates.begin() + (std::size_t)(i + 1) * num_rates);
    xmsa_emission_.assign(b.xe.begin() + (std::size_t)i * d.n_sites, b.xe.begin() + (std::size_t)(i + 1) * d.n_sites);
    vd_junction_.scaler_counts.assign(b.vds.begin() + (std::size_t)i * d.vd_sites, b.vds.begin() + (std::size_t)(i + 1) * d.vd_sites);
    if (d.has_d)
      dj_junction_.scaler_counts.assign(b.djs.begin() + (std::size_t)i * d.dj_sites, b.djs.begin() + (std::size_t)(i + 1) * d.dj_sites);
    vgerm_.scaler_count = b.gs[3 * i];
    dgerm_.scaler_count = b.gs[3 * i + 1];
    jgerm_.scaler_count = b.gs[3 * i + 2];
    loglik_cached_ = lh_loglikelihood_ = b.ll[i];
    logweight_ = lh_loglikelihood_ - rb_loglikelihood_;
    naive_sequence_ = naive_seq_samp_;
    cache_forward_ = false;
    if (last_on_device) {
      // the row's forward tables stayed on the device: score it once more as the current sample to have them (and
      // the emission and scaler counts) in the object, as the reference's object has them after its last row
      tree_ = r.tree;
      const double ll = b.ll[i];
      ScoreCurrentSample();
      loglik_cached_ = lh_loglikelihood_ = ll;
    }
  }
  if (dev_sampler) rng_ = run_gen;   // the generator stands where the serial walk would have left it
  tree_.clear();
  tree_export_cache_.clear();
  outfile.close();
  if (std::getenv("LH_TRACE"))
    std::fprintf(stderr, "LHTRACE RunPipeline busy ms: reader %.1f (%d threads)  scorer %.1f  sampler+writer %.1f\n", ms_reader,
                 reader_threads, ms_scorer, ms_sampler);
  if (!failure.empty()) {
    if (failure_is_input) throw std::invalid_argument(failure);
    throw std::runtime_error(failure);
  }
}

}  // namespace linearham
