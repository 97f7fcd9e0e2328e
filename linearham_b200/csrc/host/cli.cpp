// Command line of the B200 path: the reference's three subcommands with the reference's flag
// names (src/linearham.cpp:268-455), plus --device and --batch-size.
//
//   linearham --compute-logl --yaml-path F --cluster-ind I --hmm-param-dir D [--seed S] [--num-rates R]
//             --newick-path T --er x (six times) --pi x (four times) [--alpha A]
//   linearham --sample        (same flags) [--N n]
//   linearham --pipeline      --yaml-path F --cluster-ind I --hmm-param-dir D [--seed S] [--num-rates R]
//             --input-path revbayes.tsv --output-path out.tsv [--batch-size B]
//
// Output is what the reference prints: the log-likelihood (operator<< default precision), one
// sampled naive sequence per line, or the pipeline TSV. Errors go to stderr as "ERROR: ..." and
// the exit status is EXIT_FAILURE, as there.
#include <cstdlib>
#include <iostream>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/linearham_b200.h"
#include "phylo_hmm.hpp"

namespace {

struct Flag {
  const char* name;
  bool required, multi;
  const char* fallback;  // default when optional
  const char* help;
};

const Flag kCommon[] = {
    {"yaml-path", true, false, "", "The partis output YAML file path."},
    {"cluster-ind", true, false, "0", "An index specifying the clonal family of interest."},
    {"hmm-param-dir", true, false, "", "The directory of partis HMM germline parameter files."},
    {"seed", false, false, "0", "The RNG seed."},
    {"num-rates", false, false, "1", "The number of gamma rate categories."},
    {"device", false, false, "0", "The CUDA device ordinal."},
    {"devices", false, false, "", "CUDA devices to spread the batched work over, e.g. 0-7 or 0,2,4 (default: $LINEARHAM_DEVICES, else --device)."},
};
const Flag kTree[] = {
    {"newick-path", true, false, "", "The Newick tree file path."},
    {"er", true, true, "", "The GTR exchangeability rates."},
    {"pi", true, true, "", "The GTR stationary distribution."},
    {"alpha", false, false, "1.0", "The gamma shape parameter for among-site rate variation."},
};
const Flag kSample[] = {{"N", false, false, "1", "How many naive sequences should we sample?"}};
const Flag kPipeline[] = {
    {"input-path", true, false, "", "The RevBayes output TSV file path."},
    {"output-path", true, false, "", "The linearham output TSV file path."},
    {"batch-size", false, false, "256", "RevBayes samples scored per device call."},
};

class Args {
 public:
  template <size_t N>
  void declare(const Flag (&flags)[N]) {
    for (const Flag& f : flags) known_.push_back(&f);
  }
  void parse(const std::vector<std::string>& argv, size_t first) {
    for (size_t i = first; i < argv.size(); ++i) {
      std::string tok = argv[i], value;
      if (tok.rfind("--", 0) != 0) throw std::invalid_argument("unexpected argument '" + tok + "'");
      tok.erase(0, 2);
      bool has_value = false;
      const size_t eq = tok.find('=');
      if (eq != std::string::npos) {
        value = tok.substr(eq + 1);
        tok.erase(eq);
        has_value = true;
      }
      const Flag* f = nullptr;
      for (const Flag* k : known_)
        if (tok == k->name) f = k;
      if (!f) throw std::invalid_argument("'--" + tok + "' is not a valid argument for this subcommand");
      if (!has_value) {
        if (i + 1 >= argv.size()) throw std::invalid_argument("missing a value for '--" + tok + "'");
        value = argv[++i];
      }
      std::vector<std::string>& slot = values_[f->name];
      if (!f->multi && !slot.empty()) throw std::invalid_argument("'--" + tok + "' was given more than once");
      slot.push_back(value);
    }
    for (const Flag* k : known_)
      if (k->required && !values_.count(k->name))
        throw std::invalid_argument(std::string("required argument missing: --") + k->name);
  }
  std::string str(const char* name) const {
    auto it = values_.find(name);
    if (it != values_.end()) return it->second.front();
    for (const Flag* k : known_)
      if (name == std::string(k->name)) return k->fallback;
    return "";
  }
  int integer(const char* name) const {
    const std::string s = str(name);
    size_t used = 0;
    int v = 0;
    try {
      v = std::stoi(s, &used);
    } catch (const std::exception&) {
      used = 0;
    }
    if (used != s.size() || s.empty()) throw std::invalid_argument(std::string("--") + name + " needs an integer, got '" + s + "'");
    return v;
  }
  static double to_double(const char* name, const std::string& s) {
    size_t used = 0;
    double v = 0;
    try {
      v = std::stod(s, &used);
    } catch (const std::exception&) {
      used = 0;
    }
    if (used != s.size() || s.empty()) throw std::invalid_argument(std::string("--") + name + " needs a number, got '" + s + "'");
    return v;
  }
  double real(const char* name) const { return to_double(name, str(name)); }
  std::vector<double> reals(const char* name) const {
    std::vector<double> out;
    auto it = values_.find(name);
    if (it != values_.end())
      for (const std::string& s : it->second) out.push_back(to_double(name, s));
    return out;
  }
  void usage(std::ostream& os, const std::string& what) const {
    os << what << "\n\nArguments:\n";
    for (const Flag* k : known_)
      os << "  --" << k->name << (k->multi ? " <value> (repeat)" : " <value>") << (k->required ? "  (required)" : "")
         << "\n      " << k->help << "\n";
  }

 private:
  std::vector<const Flag*> known_;
  std::map<std::string, std::vector<std::string>> values_;
};

bool wants_help(const std::vector<std::string>& argv, size_t first) {
  for (size_t i = first; i < argv.size(); ++i)
    if (argv[i] == "-h" || argv[i] == "--help") return true;
  return false;
}

linearham::PhyloHMMPtr make_model(const Args& a) {
  auto p = std::make_shared<linearham::PhyloHMM>(a.str("yaml-path"), a.integer("cluster-ind"), a.str("hmm-param-dir"),
                                                 a.integer("seed"));
  // --devices, else LINEARHAM_DEVICES, else --device
  std::string spec = a.str("devices");
  if (spec.empty() && std::getenv("LINEARHAM_DEVICES")) spec = std::getenv("LINEARHAM_DEVICES");
  if (spec.empty()) {
    p->set_device(a.integer("device"));
  } else {
    int dev[64], n = 0;
    if (lh_parse_devices(spec.c_str(), dev, 64, &n)) throw std::invalid_argument(lh_group_last_error());
    p->set_devices(std::vector<int>(dev, dev + n));
  }
  return p;
}

}  // namespace

int main(int argc, char** argv) {
  try {
    const std::vector<std::string> args(argv, argv + argc);
    if (args.size() < 2 || args[1] == "-h" || args[1] == "--help") {
      std::cout << "A Phylo-HMM implementation for B cell receptor sequence analysis (B200 path).\n\n"
                   "Subcommands (exactly one):\n"
                   "  --compute-logl   The Phylo-HMM log-likelihood computation subcommand.\n"
                   "  --sample         The Phylo-HMM naive sequence sampling subcommand.\n"
                   "  --pipeline       The linearham pipeline subcommand.\n"
                   "Use '<subcommand> --help' for its arguments.\n";
      return args.size() < 2 ? EXIT_FAILURE : EXIT_SUCCESS;
    }
    const std::string sub = args[1];
    Args a;
    a.declare(kCommon);
    if (sub == "--compute-logl" || sub == "--sample") {
      a.declare(kTree);
      if (sub == "--sample") a.declare(kSample);
      if (wants_help(args, 2)) {
        a.usage(std::cout, sub == "--sample" ? "The Phylo-HMM naive sequence sampler." : "The Phylo-HMM log-likelihood calculator.");
        return EXIT_SUCCESS;
      }
      a.parse(args, 2);
      linearham::PhyloHMMPtr model = make_model(a);
      model->InitializePhyloParameters(a.str("newick-path"), a.reals("er"), a.reals("pi"), a.real("alpha"),
                                       a.integer("num-rates"));
      model->InitializePhyloEmission();
      if (sub == "--compute-logl") {
        std::cout << model->LogLikelihood() << std::endl;
      } else {
        const int n = a.integer("N");
        for (int i = 0; i < n; ++i) std::cout << model->SampleNaiveSequence() << std::endl;
      }
      return EXIT_SUCCESS;
    }
    if (sub == "--pipeline") {
      a.declare(kPipeline);
      if (wants_help(args, 2)) {
        a.usage(std::cout, "The linearham pipeline.");
        return EXIT_SUCCESS;
      }
      a.parse(args, 2);
      const int batch = a.integer("batch-size");
      if (batch < 1) throw std::invalid_argument("--batch-size must be positive");
      make_model(a)->RunPipeline(a.str("input-path"), a.str("output-path"), a.integer("num-rates"), batch);
      return EXIT_SUCCESS;
    }
    throw std::invalid_argument("'" + sub + "' is not a valid subcommand.");
  } catch (const std::exception& e) {
    std::cerr << "ERROR: " << e.what() << std::endl;
  }
  return EXIT_FAILURE;
}
