// cfg (INI) reader with the reference's semantics: sections in file order, key=value pairs,
// '#' / ';' / blank lines skipped, whitespace stripped (parser.cpp:59-100, option_list.cpp:134-243),
// typed lookups with defaults (FindOptionInt/Float/Str(Quiet), option_list.h:35-42).
#pragma once
#include <stdlib.h>
#include <string.h>

#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace yc {

struct Section {
  std::string name;
  std::vector<std::pair<std::string, std::string>> kv;
  mutable std::map<std::string, bool> used;

  const char* find(const char* key) const {
    // like option_find: the LAST occurrence wins is not the reference's rule; it returns the first
    for (auto& p : kv)
      if (p.first == key) { used[key] = true; return p.second.c_str(); }
    return nullptr;
  }
  int get_int(const char* key, int def) const { const char* v = find(key); return v ? atoi(v) : def; }
  float get_float(const char* key, float def) const { const char* v = find(key); return v ? (float)atof(v) : def; }
  std::string get_str(const char* key, const char* def) const { const char* v = find(key); return v ? v : (def ? def : ""); }
  bool has(const char* key) const { return find(key) != nullptr; }
};

static inline std::string strip_ws(const std::string& s) {
  std::string o;
  for (char c : s)
    if (c != ' ' && c != '\t' && c != '\r' && c != '\n') o.push_back(c);   // utils.cpp strip()
  return o;
}

// returns false + message on malformed input
static inline bool parse_cfg_text(const std::string& text, std::vector<Section>& out, std::string& err) {
  std::istringstream in(text);
  std::string raw;
  int lineno = 0;
  while (std::getline(in, raw)) {
    ++lineno;
    std::string line = strip_ws(raw);
    if (line.empty()) continue;
    switch (line[0]) {
      case '[': {
        Section s;
        size_t e = line.find(']');
        s.name = line.substr(1, e == std::string::npos ? std::string::npos : e - 1);
        out.push_back(s);
        break;
      }
      case '#':
      case ';':
        break;
      default: {
        size_t eq = line.find('=');
        if (eq == std::string::npos || out.empty()) {
          err = "cfg line " + std::to_string(lineno) + ": expected key=value: " + raw;
          return false;
        }
        out.back().kv.emplace_back(line.substr(0, eq), line.substr(eq + 1));
      }
    }
  }
  if (out.empty() || !(out[0].name == "net" || out[0].name == "network")) {
    err = "First section must be [net] or [network]";   // parser.cpp:1091-1092
    return false;
  }
  return true;
}

static inline std::vector<int> parse_int_list(const std::string& s) {
  std::vector<int> v;
  std::stringstream ss(s);
  std::string tok;
  while (std::getline(ss, tok, ','))
    if (!tok.empty()) v.push_back(atoi(tok.c_str()));
  return v;
}
static inline std::vector<float> parse_float_list(const std::string& s) {
  std::vector<float> v;
  std::stringstream ss(s);
  std::string tok;
  while (std::getline(ss, tok, ','))
    if (!tok.empty()) v.push_back((float)atof(tok.c_str()));
  return v;
}

}  // namespace yc
