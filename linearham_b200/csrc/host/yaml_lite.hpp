// Minimal YAML reader for the two file kinds the phylo-HMM path ingests:
// partis HMM parameter files (block style with `!!python/object` tags, inline flow maps/lists,
// comments) and partis cluster files (JSON-style flow collections spanning lines).
// The reference uses yaml-cpp 0.5.3 (src/HMM.cpp:30, src/VDJGermline.cpp:70); that library is
// not available here, and only this subset is needed: block and flow maps / sequences, plain and
// quoted scalars, `null`, tags (ignored), `#` comments. Maps keep document order.
#pragma once
#include <string>
#include <utility>
#include <vector>

namespace lhb {
namespace yaml {

struct Node {
  enum Kind { Null, Scalar, Map, Seq };
  Kind kind = Null;
  std::string scalar;
  std::vector<std::pair<std::string, Node>> map;
  std::vector<Node> seq;

  bool is_null() const { return kind == Null; }
  bool is_map() const { return kind == Map; }
  bool is_seq() const { return kind == Seq; }
  std::size_t size() const { return kind == Map ? map.size() : kind == Seq ? seq.size() : 0; }
  const Node* find(const std::string& key) const;
  const Node& at(const std::string& key) const;   // throws std::runtime_error when missing
  const Node& at(std::size_t i) const;            // sequence element, throws when out of range
  const std::string& as_string() const;           // throws unless scalar
  double as_double() const;
  int as_int() const;
  bool as_bool() const;
};

Node load_file(const std::string& path);  // throws std::runtime_error (unreadable / malformed)
Node load_string(const std::string& text);

}  // namespace yaml
}  // namespace lhb
