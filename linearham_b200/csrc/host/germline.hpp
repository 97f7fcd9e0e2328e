// Per-gene HMM parameters read from a partis HMM parameter directory: what the reference keeps in
// its Germline / NTInsertion / NPadding / VDJGermline classes (src/Germline.cpp:20-115,
// src/NTInsertion.cpp:21-104, src/NPadding.cpp:22-109, src/VDJGermline.cpp:46-108), here one
// plain struct per gene. Parsed once per family on the host.
#pragma once
#include <map>
#include <string>
#include <vector>

namespace lhb {

struct GermlineGene {
  char type = 'V';               // 'V', 'D' or 'J' (GermlineType)
  std::string name;              // with '*' and '/' restored (IGHV1-2*01)
  std::string alphabet;          // sorted nucleotide alphabet, "ACGT"
  double gene_prob = 0.0;
  int length() const { return (int)bases.size(); }
  // germline-encoded states (Germline)
  std::vector<double> landing_in, landing_out;  // [n]
  std::vector<double> transition;               // [n-1] to the next germline position
  std::vector<double> emission;                 // [4][n] column-major (unused on the phylo path)
  std::vector<int> bases;                       // [n] germline base indices
  // non-templated insertions to the left of D and J genes (NTInsertion), 4 = alphabet size
  std::vector<double> nti_landing_in;           // [4]
  std::vector<double> nti_landing_out;          // [4][n] column-major (base, germline position)
  std::vector<double> nti_transition;           // [4][4] column-major (from, to)
  std::vector<double> nti_emission;             // [4][4] column-major (emitted, state)
  // N padding left of V / right of J (NPadding)
  double n_transition = 0.0;
  std::vector<double> n_emission;               // [4]
};

// (gene name -> parameters) for every IG[HKL][VDJ]*_star_*.yaml in the directory; IGK/IGL "D"
// files are skipped as in the reference. Throws std::runtime_error for a missing directory or a
// file that violates the layout the reference asserts.
std::map<std::string, GermlineGene> load_germline_dir(const std::string& hmm_param_dir);

}  // namespace lhb
