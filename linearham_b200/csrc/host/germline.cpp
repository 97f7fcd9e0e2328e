#include "germline.hpp"

#include <dirent.h>

#include <algorithm>
#include <cmath>
#include <stdexcept>

#include "yaml_lite.hpp"

namespace lhb {
namespace {

constexpr double kEps = 1e-6;  // src/utils.hpp:20

void require(bool ok, const std::string& gene, const char* what) {
  if (!ok) throw std::runtime_error("HMM parameter file of " + gene + ": " + what);
}

std::string replace_all(std::string s, const std::string& from, const std::string& to) {
  for (std::size_t p = 0; (p = s.find(from, p)) != std::string::npos; p += to.size()) s.replace(p, from.size(), to);
  return s;
}

// "<gene>_<digits>" -> position, else -1   (GetGermlineStateRegex, src/utils.cpp:75-80)
int germline_position(const std::string& state, const std::string& gene) {
  if (state.size() <= gene.size() + 1 || state.compare(0, gene.size(), gene) != 0 || state[gene.size()] != '_') return -1;
  int v = 0;
  for (std::size_t i = gene.size() + 1; i < state.size(); ++i) {
    if (state[i] < '0' || state[i] > '9') return -1;
    v = v * 10 + (state[i] - '0');
  }
  return v;
}

// "insert_left_<base>" with base in the alphabet -> base index, else -1 (GetNTIStateRegex)
int nti_base(const std::string& state, const std::string& alphabet) {
  static const std::string pre = "insert_left_";
  if (state.size() != pre.size() + 1 || state.compare(0, pre.size(), pre) != 0) return -1;
  const std::size_t p = alphabet.find(state.back());
  return p == std::string::npos ? -1 : (int)p;
}

int base_index(const std::string& alphabet, char c, const std::string& gene) {
  const std::size_t p = alphabet.find(c);
  require(p != std::string::npos, gene, "base outside the alphabet");
  return (int)p;
}

// (state, probability) pairs of a YAML map, probabilities summing to 1 (ParseStringProbMap)
std::vector<std::pair<std::string, double>> prob_map(const yaml::Node& n, const std::string& gene) {
  require(n.is_map(), gene, "expected a map of probabilities");
  std::vector<std::pair<std::string, double>> out;
  double sum = 0;
  for (const auto& kv : n.map) {
    out.emplace_back(kv.first, kv.second.as_double());
    sum += out.back().second;
  }
  require(std::fabs(sum - 1) <= kEps, gene, "probabilities do not sum to 1");
  return out;
}

GermlineGene parse_gene(const yaml::Node& root, char type, const std::string& file_gene) {
  GermlineGene g;
  g.type = type;
  // alphabet (GetAlphabet, src/utils.cpp:40-48)
  for (const auto& c : root.at("tracks").at("nukes").seq) g.alphabet.push_back(c.as_string().at(0));
  std::sort(g.alphabet.begin(), g.alphabet.end());
  const int A = (int)g.alphabet.size();
  const std::string raw_name = root.at("name").as_string();
  g.name = replace_all(replace_all(raw_name, "_star_", "*"), "_slash_", "/");
  g.gene_prob = root.at("extras").at("gene_prob").as_double();
  const yaml::Node& states = root.at("states");
  const int n_states = (int)states.size();

  // first / last state whose name contains the gene name (FindGermlineStartEnd)
  int gstart = 0, gend = n_states - 1;
  while (gstart < n_states && states.at(gstart).at("name").as_string().find(raw_name) == std::string::npos) ++gstart;
  while (gend >= 0 && states.at(gend).at("name").as_string().find(raw_name) == std::string::npos) --gend;
  require(gstart <= gend, file_gene, "no germline states");
  require(gstart == 2 || gstart == A + 1, file_gene, "unexpected number of insert_left states");
  require(gend == n_states - 1 || gend == n_states - 2, file_gene, "unexpected states after the germline states");
  const int n = gend - gstart + 1;
  g.landing_in.assign(n, 0.0);
  g.landing_out.assign(n, 0.0);
  g.transition.assign(n > 0 ? n - 1 : 0, 0.0);
  g.emission.assign((std::size_t)A * n, 0.0);
  g.bases.assign(n, 0);

  const yaml::Node& init = states.at(0);
  require(init.at("name").as_string() == "init", file_gene, "first state is not 'init'");
  for (const auto& sp : prob_map(init.at("transitions"), file_gene)) {
    const int pos = germline_position(sp.first, raw_name);
    if (pos >= 0) {
      require(pos < n, file_gene, "landing-in position out of range");
      g.landing_in[pos] = sp.second;
    } else {
      require(sp.first.find("insert_left_") != std::string::npos, file_gene, "unexpected init transition");
    }
  }

  for (int i = gstart; i <= gend; ++i) {
    const yaml::Node& st = states.at(i);
    const int gi = germline_position(st.at("name").as_string(), raw_name);
    require(gi == i - gstart, file_gene, "germline states out of order");
    for (const auto& sp : prob_map(st.at("transitions"), file_gene)) {
      const int pos = germline_position(sp.first, raw_name);
      if (pos >= 0) {
        require(pos == gi + 1, file_gene, "germline state does not continue to the next position");
        g.transition.at(gi) = sp.second;
      } else if (sp.first == "end") {
        g.landing_out[gi] = sp.second;
      } else {
        require(sp.first == "insert_right_N", file_gene, "unexpected germline transition");
      }
    }
    const yaml::Node& em = st.at("emissions");
    require(em.at("track").as_string() == "nukes", file_gene, "emission track is not 'nukes'");
    for (const auto& sp : prob_map(em.at("probs"), file_gene))
      g.emission[(std::size_t)gi * A + base_index(g.alphabet, sp.first.at(0), file_gene)] = sp.second;
    g.bases[gi] = base_index(g.alphabet, st.at("extras").at("germline").as_string().at(0), file_gene);
  }

  if (type == 'D' || type == 'J') {  // NTInsertion
    require(gstart == A + 1, file_gene, "D/J gene without the insert_left_[ACGT] states");
    g.nti_landing_in.assign(A, 0.0);
    g.nti_landing_out.assign((std::size_t)A * n, 0.0);
    g.nti_transition.assign((std::size_t)A * A, 0.0);
    g.nti_emission.assign((std::size_t)A * A, 0.0);
    for (const auto& sp : prob_map(init.at("transitions"), file_gene)) {
      const int b = nti_base(sp.first, g.alphabet);
      if (b >= 0) g.nti_landing_in[b] = sp.second;
      else require(germline_position(sp.first, raw_name) >= 0, file_gene, "unexpected init transition");
    }
    for (int i = 1; i <= A; ++i) {
      const yaml::Node& st = states.at(i);
      const int b = nti_base(st.at("name").as_string(), g.alphabet);
      require(b >= 0, file_gene, "expected an insert_left_[ACGT] state");
      for (const auto& sp : prob_map(st.at("transitions"), file_gene)) {
        const int pos = germline_position(sp.first, raw_name);
        if (pos >= 0) {
          require(pos < n, file_gene, "NTI landing-out position out of range");
          g.nti_landing_out[(std::size_t)pos * A + b] = sp.second;
        } else {
          const int nb = nti_base(sp.first, g.alphabet);
          require(nb >= 0, file_gene, "unexpected NTI transition");
          g.nti_transition[(std::size_t)nb * A + b] = sp.second;
        }
      }
      const yaml::Node& em = st.at("emissions");
      require(em.at("track").as_string() == "nukes", file_gene, "emission track is not 'nukes'");
      for (const auto& sp : prob_map(em.at("probs"), file_gene))
        g.nti_emission[(std::size_t)b * A + base_index(g.alphabet, sp.first.at(0), file_gene)] = sp.second;
    }
  }

  if (type == 'V' || type == 'J') {  // NPadding
    int n_index, check_index;
    std::string n_name, next_name;
    if (gstart == 2) {
      n_index = gstart - 1;
      check_index = gstart - 2;
      n_name = "insert_left_N";
      next_name = raw_name + "_0";
    } else {
      require(gend == n_states - 2, file_gene, "no insert_left_N / insert_right_N state");
      n_index = gend + 1;
      check_index = gend;
      n_name = "insert_right_N";
      next_name = "end";
    }
    const yaml::Node& ns = states.at(n_index);
    require(ns.at("name").as_string() == n_name, file_gene, "N-padding state has an unexpected name");
    // the padding state and the state before it must agree (sorted by name, as std::map does)
    std::map<std::string, double> a, b;
    for (const auto& kv : ns.at("transitions").map) a[kv.first] = kv.second.as_double();
    for (const auto& kv : states.at(check_index).at("transitions").map) b[kv.first] = kv.second.as_double();
    require(a.size() == b.size(), file_gene, "N-padding transitions differ from the preceding state");
    auto ia = a.begin();
    auto ib = b.begin();
    for (; ia != a.end(); ++ia, ++ib) {
      require(ia->first == ib->first && std::fabs(ia->second - ib->second) <= kEps, file_gene,
              "N-padding transitions differ from the preceding state");
      if (ia->first == n_name) g.n_transition = ia->second;
      else require(ia->first == next_name, file_gene, "unexpected N-padding transition");
    }
    g.n_emission.assign(A, 0.0);
    const yaml::Node& em = ns.at("emissions");
    require(em.at("track").as_string() == "nukes", file_gene, "emission track is not 'nukes'");
    for (const auto& sp : prob_map(em.at("probs"), file_gene)) {
      require(sp.second == 0.25, file_gene, "N-padding emission is not 0.25");
      g.n_emission[base_index(g.alphabet, sp.first.at(0), file_gene)] = sp.second;
    }
    require(ns.at("extras").at("germline").as_string() == "N", file_gene, "N-padding germline base is not N");
    require(ns.at("extras").at("ambiguous_emission_prob").as_double() == 0.25, file_gene,
            "N-padding ambiguous_emission_prob is not 0.25");
  }
  return g;
}

}  // namespace

std::map<std::string, GermlineGene> load_germline_dir(const std::string& dir_in) {
  std::string dir = dir_in;
  if (dir.empty() || dir.back() != '/') dir += "/";
  DIR* d = opendir(dir.c_str());
  if (!d) throw std::runtime_error("--hmm-param-dir \"" + dir + "\" does not exist");
  std::map<std::string, GermlineGene> genes;
  std::vector<std::string> files;
  while (dirent* e = readdir(d)) files.push_back(e->d_name);
  closedir(d);
  std::sort(files.begin(), files.end());
  for (const std::string& fn : files) {
    // ^(IG([HKL])([VDJ]).*_star_.*)\.yaml$   (src/VDJGermline.cpp:56)
    if (fn.size() < 10 || fn.compare(0, 2, "IG") != 0 || fn.compare(fn.size() - 5, 5, ".yaml") != 0) continue;
    const char locus = fn[2], type = fn[3];
    if ((locus != 'H' && locus != 'K' && locus != 'L') || (type != 'V' && type != 'D' && type != 'J')) continue;
    if (fn.find("_star_", 4) == std::string::npos) continue;
    if (type == 'D' && (locus == 'K' || locus == 'L')) continue;
    const std::string stem = fn.substr(0, fn.size() - 5);
    GermlineGene g = parse_gene(yaml::load_file(dir + fn), type, stem);
    const std::string key = replace_all(replace_all(stem, "_star_", "*"), "_slash_", "/");
    genes.emplace(key, std::move(g));
  }
  if (genes.empty()) throw std::runtime_error("no germline HMM parameter files in \"" + dir + "\"");
  for (const auto& kv : genes)
    if (kv.second.alphabet != genes.begin()->second.alphabet)
      throw std::runtime_error("germline genes disagree on the nucleotide alphabet");
  return genes;
}

}  // namespace lhb
