#include "yaml_lite.hpp"

#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace lhb {
namespace yaml {

namespace {

[[noreturn]] void bad(const std::string& what) { throw std::runtime_error("yaml: " + what); }

struct Line {
  int indent;
  std::string text;  // without indentation, comments and trailing blanks
};

std::string rtrim(std::string s) {
  while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r' || s.back() == '\n')) s.pop_back();
  return s;
}

// removes a trailing "# comment" that is not inside quotes
std::string strip_comment(const std::string& s) {
  char quote = 0;
  for (std::size_t i = 0; i < s.size(); ++i) {
    const char c = s[i];
    if (quote) {
      if (c == quote) quote = 0;
    } else if (c == '"' || c == '\'') {
      quote = c;
    } else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) {
      return s.substr(0, i);
    }
  }
  return s;
}

std::vector<Line> split_lines(const std::string& text) {
  std::vector<Line> out;
  std::istringstream in(text);
  std::string raw;
  while (std::getline(in, raw)) {
    std::string s = rtrim(strip_comment(raw));
    std::size_t ind = 0;
    while (ind < s.size() && s[ind] == ' ') ++ind;
    s = s.substr(ind);
    if (s.empty() || s == "---" || s == "...") continue;
    out.push_back({(int)ind, s});
  }
  return out;
}

std::string unquote(const std::string& s) {
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\'')))
    return s.substr(1, s.size() - 2);
  return s;
}

Node make_scalar(const std::string& raw) {
  Node n;
  const std::string t = rtrim(raw);
  const bool quoted = t.size() >= 2 && (t.front() == '"' || t.front() == '\'');
  if (!quoted && (t.empty() || t == "null" || t == "~" || t == "Null" || t == "NULL")) return n;
  n.kind = Node::Scalar;
  n.scalar = unquote(t);
  return n;
}

// drops leading "!tag" tokens
std::string strip_tags(std::string s) {
  for (;;) {
    std::size_t i = 0;
    while (i < s.size() && s[i] == ' ') ++i;
    s = s.substr(i);
    if (s.empty() || s[0] != '!') return s;
    std::size_t e = s.find(' ');
    s = (e == std::string::npos) ? std::string() : s.substr(e + 1);
  }
}

// ---- flow style ------------------------------------------------------------------------

struct Flow {
  const std::string& s;
  std::size_t p = 0;
  explicit Flow(const std::string& str) : s(str) {}
  void ws() {
    while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) ++p;
  }
  std::string token(bool is_key) {
    ws();
    if (p >= s.size()) bad("unexpected end of flow collection");
    if (s[p] == '"' || s[p] == '\'') {
      const char q = s[p];
      std::size_t e = p + 1;
      std::string out;
      while (e < s.size() && s[e] != q) {
        if (s[e] == '\\' && q == '"' && e + 1 < s.size()) ++e;
        out.push_back(s[e]);
        ++e;
      }
      if (e >= s.size()) bad("unterminated quoted string");
      p = e + 1;
      return std::string(1, q) + out + std::string(1, q);
    }
    std::size_t e = p;
    while (e < s.size()) {
      const char c = s[e];
      if (c == ',' || c == '}' || c == ']' || c == '{' || c == '[') break;
      if (is_key && c == ':' && (e + 1 >= s.size() || s[e + 1] == ' ' || s[e + 1] == '\n')) break;
      ++e;
    }
    std::string out = rtrim(s.substr(p, e - p));
    p = e;
    return out;
  }
  Node value() {
    ws();
    if (p >= s.size()) bad("unexpected end of flow collection");
    if (s[p] == '{') {
      ++p;
      Node n;
      n.kind = Node::Map;
      for (;;) {
        ws();
        if (p < s.size() && s[p] == '}') { ++p; break; }
        std::string key = unquote(token(true));
        ws();
        if (p >= s.size() || s[p] != ':') bad("expected ':' after key '" + key + "'");
        ++p;
        n.map.emplace_back(key, value());
        ws();
        if (p < s.size() && s[p] == ',') { ++p; continue; }
        if (p < s.size() && s[p] == '}') { ++p; break; }
        bad("expected ',' or '}' in flow map");
      }
      return n;
    }
    if (s[p] == '[') {
      ++p;
      Node n;
      n.kind = Node::Seq;
      for (;;) {
        ws();
        if (p < s.size() && s[p] == ']') { ++p; break; }
        n.seq.push_back(value());
        ws();
        if (p < s.size() && s[p] == ',') { ++p; continue; }
        if (p < s.size() && s[p] == ']') { ++p; break; }
        bad("expected ',' or ']' in flow sequence");
      }
      return n;
    }
    return make_scalar(token(false));
  }
};

int bracket_balance(const std::string& s) {
  int depth = 0;
  char quote = 0;
  for (char c : s) {
    if (quote) {
      if (c == quote) quote = 0;
    } else if (c == '"' || c == '\'') {
      quote = c;
    } else if (c == '{' || c == '[') {
      ++depth;
    } else if (c == '}' || c == ']') {
      --depth;
    }
  }
  return depth;
}

// ---- block style -----------------------------------------------------------------------

struct Block {
  std::vector<Line> lines;
  std::size_t i = 0;

  // splits "key: rest" / "key:"; returns false when the text is not a map entry
  static bool split_key(const std::string& t, std::string* key, std::string* rest) {
    char quote = 0;
    for (std::size_t k = 0; k < t.size(); ++k) {
      const char c = t[k];
      if (quote) {
        if (c == quote) quote = 0;
      } else if (c == '"' || c == '\'') {
        quote = c;
      } else if (c == '{' || c == '[') {
        return false;
      } else if (c == ':' && (k + 1 == t.size() || t[k + 1] == ' ')) {
        *key = unquote(rtrim(t.substr(0, k)));
        *rest = (k + 1 < t.size()) ? t.substr(k + 2) : std::string();
        return true;
      }
    }
    return false;
  }

  // a flow collection that starts in `first` may continue on the following lines
  Node flow_from(std::string first) {
    while (bracket_balance(first) > 0) {
      if (i >= lines.size()) bad("unterminated flow collection");
      first += "\n" + lines[i].text;
      ++i;
    }
    Flow f(first);
    Node n = f.value();
    return n;
  }

  // value that follows "key:" or "- " on the current line (already consumed), `ind` = indent of
  // the owning line
  Node inline_or_nested(const std::string& rest_raw, int ind, bool owner_is_map) {
    const std::string rest = strip_tags(rest_raw);
    if (!rest.empty()) {
      if (rest[0] == '{' || rest[0] == '[') return flow_from(rest);
      return make_scalar(rest);
    }
    if (i >= lines.size()) return Node();
    if (lines[i].indent > ind) return node(lines[i].indent);
    // a block sequence may sit at the same indentation as the key that owns it
    if (owner_is_map && lines[i].indent == ind && (lines[i].text == "-" || lines[i].text.rfind("- ", 0) == 0))
      return node(ind);
    return Node();
  }

  Node node(int ind) {
    if (i >= lines.size() || lines[i].indent < ind) return Node();
    const std::string& t = lines[i].text;
    if (t == "-" || t.rfind("- ", 0) == 0) return sequence(lines[i].indent);
    std::string key, rest;
    if (split_key(t, &key, &rest)) return mapping(lines[i].indent);
    if (t[0] == '{' || t[0] == '[') {
      std::string first = t;
      ++i;
      return flow_from(first);
    }
    Node n = make_scalar(strip_tags(t));
    ++i;
    return n;
  }

  Node mapping(int ind) {
    Node n;
    n.kind = Node::Map;
    while (i < lines.size() && lines[i].indent == ind) {
      std::string key, rest;
      if (!split_key(lines[i].text, &key, &rest)) break;
      ++i;
      n.map.emplace_back(key, inline_or_nested(rest, ind, true));
    }
    return n;
  }

  Node sequence(int ind) {
    Node n;
    n.kind = Node::Seq;
    while (i < lines.size() && lines[i].indent == ind && (lines[i].text == "-" || lines[i].text.rfind("- ", 0) == 0)) {
      std::string item = lines[i].text.size() > 2 ? strip_tags(lines[i].text.substr(2)) : std::string();
      std::string key, rest;
      if (!item.empty() && split_key(item, &key, &rest)) {
        // "- key: value" opens a map whose other keys are indented by two more columns
        lines[i].indent = ind + 2;
        lines[i].text = item;
        n.seq.push_back(mapping(ind + 2));
      } else {
        ++i;
        n.seq.push_back(inline_or_nested(item, ind, false));
      }
    }
    return n;
  }
};

}  // namespace

const Node* Node::find(const std::string& key) const {
  if (kind != Map) return nullptr;
  for (const auto& kv : map)
    if (kv.first == key) return &kv.second;
  return nullptr;
}
const Node& Node::at(const std::string& key) const {
  const Node* n = find(key);
  if (!n) bad("missing key '" + key + "'");
  return *n;
}
const Node& Node::at(std::size_t idx) const {
  if (kind != Seq || idx >= seq.size()) bad("sequence index " + std::to_string(idx) + " out of range");
  return seq[idx];
}
const std::string& Node::as_string() const {
  if (kind != Scalar) bad("expected a scalar");
  return scalar;
}
double Node::as_double() const {
  const std::string& s = as_string();
  char* end = nullptr;
  const double v = std::strtod(s.c_str(), &end);
  if (end == s.c_str() || *end != '\0') bad("'" + s + "' is not a number");
  return v;
}
int Node::as_int() const {
  const std::string& s = as_string();
  char* end = nullptr;
  const long v = std::strtol(s.c_str(), &end, 10);
  if (end == s.c_str() || *end != '\0') bad("'" + s + "' is not an integer");
  return (int)v;
}
bool Node::as_bool() const {
  const std::string& s = as_string();
  if (s == "true" || s == "True" || s == "TRUE" || s == "yes") return true;
  if (s == "false" || s == "False" || s == "FALSE" || s == "no") return false;
  bad("'" + s + "' is not a boolean");
}

Node load_string(const std::string& text) {
  Block b;
  b.lines = split_lines(text);
  // a document may open with a bare tag line (!!python/object:...)
  while (b.i < b.lines.size() && b.lines[b.i].text[0] == '!' && b.lines[b.i].text.find(' ') == std::string::npos) ++b.i;
  if (b.i >= b.lines.size()) return Node();
  return b.node(b.lines[b.i].indent);
}

Node load_file(const std::string& path) {
  std::ifstream in(path);
  if (!in) throw std::runtime_error("cannot open " + path);
  std::stringstream ss;
  ss << in.rdbuf();
  try {
    return load_string(ss.str());
  } catch (const std::runtime_error& e) {
    throw std::runtime_error(path + ": " + e.what());
  }
}

}  // namespace yaml
}  // namespace lhb
