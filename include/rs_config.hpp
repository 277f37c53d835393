// rs_config.hpp -- the worker's configuration file, read as the reference reads it.
//
// The reference parses `service.cfg` with libconfig 1.5 and looks every setting up as a string
// (src/service/service.cpp:1386-1442: cfg.readFile, cfg.lookup("size_of_sample").c_str() -> atoi, ...,
// cfg.lookup("rocksdb") as an array of strings; template: demo/TEMPLATE.service.cfg).  This header parses that file
// format without the library: `name = "string";` settings and `name = [ "a", "b" ];` arrays (also ( ) lists), with
// libconfig's comments (`//`, `#`, `/* */`), optional `;` or `,` terminators, adjacent string concatenation and the
// escapes \\ \" \n \r \t \f \xHH (an unknown escape is kept as written, as libconfig keeps it).  Scalars that are
// not strings (numbers, booleans) and groups are accepted and kept
// as "not a string": looking one up as a string fails, as the reference's c_str() on such a setting throws.
// tests/test_service_shell.py compares it with the reference's own vendored libconfig on the same text.
#ifndef RS_CONFIG_HPP
#define RS_CONFIG_HPP

#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace rsb200 {

struct ConfigError : std::runtime_error {
  explicit ConfigError(const std::string& m) : std::runtime_error(m) {}
};

class Config {
 public:
  static Config parse(const std::string& text) {
    Config c;
    Parser p{text, 0, 1};
    p.settings(c, /*in_group=*/false);
    return c;
  }
  static Config read_file(const std::string& path) {
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f) throw ConfigError("cannot open " + path);  // FileIOException in the reference
    std::stringstream ss;
    ss << f.rdbuf();
    return parse(ss.str());
  }

  bool exists(const std::string& name) const { return kinds_.count(name) != 0; }
  // cfg.lookup(name).c_str(): throws when absent (SettingNotFoundException) or not a string (SettingTypeException)
  const std::string& lookup_string(const std::string& name) const {
    std::map<std::string, std::string>::const_iterator it = strings_.find(name);
    if (it == strings_.end()) throw ConfigError(exists(name) ? "setting is not a string: " + name : "setting not found: " + name);
    return it->second;
  }
  const std::vector<std::string>& lookup_array(const std::string& name) const {
    std::map<std::string, std::vector<std::string> >::const_iterator it = arrays_.find(name);
    if (it == arrays_.end()) throw ConfigError(exists(name) ? "setting is not a list of strings: " + name : "setting not found: " + name);
    return it->second;
  }

 private:
  std::map<std::string, std::string> strings_;
  std::map<std::string, std::vector<std::string> > arrays_;
  std::map<std::string, int> kinds_;  // 0 string, 1 array of strings, 2 anything else

  struct Parser {
    const std::string& s;
    size_t i;
    int line;

    [[noreturn]] void fail(const std::string& what) const {
      std::ostringstream o;
      o << "config line " << line << ": " << what;
      throw ConfigError(o.str());
    }
    void skip() {  // white space and the three comment forms
      for (;;) {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n' || s[i] == '\f')) {
          if (s[i] == '\n') ++line;
          ++i;
        }
        if (i < s.size() && s[i] == '#') {
          while (i < s.size() && s[i] != '\n') ++i;
        } else if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '/') {
          while (i < s.size() && s[i] != '\n') ++i;
        } else if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '*') {
          i += 2;
          while (i + 1 < s.size() && !(s[i] == '*' && s[i + 1] == '/')) {
            if (s[i] == '\n') ++line;
            ++i;
          }
          i = i + 1 >= s.size() ? s.size() : i + 2;  // libconfig reads an unterminated comment to the end of the file
        } else {
          return;
        }
      }
    }
    static bool name_char(char c, bool first) {
      return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || c == '*' || c == '_' ||
             (!first && ((c >= '0' && c <= '9') || c == '-'));
    }
    std::string name() {
      const size_t b = i;
      while (i < s.size() && name_char(s[i], i == b)) ++i;
      if (i == b) fail("setting name expected");
      return s.substr(b, i - b);
    }
    std::string one_string() {  // at the opening quote
      std::string out;
      ++i;
      for (;;) {
        if (i >= s.size()) fail("unterminated string");
        const char c = s[i++];
        if (c == '"') return out;
        if (c == '\n') ++line;
        if (c != '\\') {
          out.push_back(c);
          continue;
        }
        if (i >= s.size()) fail("unterminated string");
        const char e = s[i++];
        switch (e) {
          case 'n': out.push_back('\n'); break;
          case 'r': out.push_back('\r'); break;
          case 't': out.push_back('\t'); break;
          case 'f': out.push_back('\f'); break;
          case '\\': out.push_back('\\'); break;
          case '"': out.push_back('"'); break;
          case 'x': {
            if (i + 1 >= s.size()) fail("bad \\x escape");
            out.push_back((char)strtol(s.substr(i, 2).c_str(), 0, 16));
            i += 2;
            break;
          }
          default:  // libconfig keeps an escape it does not know as written
            out.push_back('\\');
            out.push_back(e);
        }
      }
    }
    std::string string_value() {  // adjacent strings concatenate
      std::string out = one_string();
      for (;;) {
        skip();
        if (i < s.size() && s[i] == '"') out += one_string();
        else return out;
      }
    }
    void other_scalar() {  // number / boolean / anything up to a delimiter
      const size_t b = i;
      while (i < s.size() && s[i] != ';' && s[i] != ',' && s[i] != '\n' && s[i] != ']' && s[i] != ')' && s[i] != '}' &&
             s[i] != ' ' && s[i] != '\t' && s[i] != '\r')
        ++i;
      if (i == b) fail("value expected");
    }
    // [ ... ] or ( ... ): returns true when every element is a string
    bool sequence(char close, std::vector<std::string>& out) {
      bool all_strings = true;
      ++i;
      for (;;) {
        skip();
        if (i >= s.size()) fail("unterminated list");
        if (s[i] == close) {
          ++i;
          return all_strings;
        }
        if (s[i] == '"') out.push_back(string_value());
        else if (s[i] == '[' || s[i] == '(') {
          std::vector<std::string> inner;
          sequence(s[i] == '[' ? ']' : ')', inner);
          all_strings = false;
        } else if (s[i] == '{') {
          Config inner;
          ++i;
          settings(inner, true);
          all_strings = false;
        } else {
          other_scalar();
          all_strings = false;
        }
        skip();
        if (i < s.size() && s[i] == ',') ++i;
      }
    }
    void settings(Config& c, bool in_group) {
      for (;;) {
        skip();
        if (i >= s.size()) {
          if (in_group) fail("unterminated group");
          return;
        }
        if (in_group && s[i] == '}') {
          ++i;
          return;
        }
        const std::string key = name();
        skip();
        if (i >= s.size() || (s[i] != '=' && s[i] != ':')) fail("'=' expected after " + key);
        ++i;
        skip();
        if (i >= s.size()) fail("value expected for " + key);
        if (c.kinds_.count(key)) fail("duplicate setting " + key);
        if (s[i] == '"') {
          c.strings_[key] = string_value();
          c.kinds_[key] = 0;
        } else if (s[i] == '[' || s[i] == '(') {
          std::vector<std::string> v;
          const bool ok = sequence(s[i] == '[' ? ']' : ')', v);
          if (ok) c.arrays_[key] = v;
          c.kinds_[key] = ok ? 1 : 2;
        } else if (s[i] == '{') {
          Config inner;
          ++i;
          settings(inner, true);
          c.kinds_[key] = 2;
        } else {
          other_scalar();
          c.kinds_[key] = 2;
        }
        skip();
        if (i < s.size() && (s[i] == ';' || s[i] == ',')) ++i;
      }
    }
  };
};

// The worker's settings as main() of the reference reads them (service.cpp:1386-1442): the four optional ones keep
// the reference's defaults (service.cpp:56-60), the others are required.
struct WorkerConfig {
  int size_of_sample = 2;            // sizeofSample, service.cpp:59
  bool has_other_meta_data = false;  // service.cpp:60
  size_t max_read_length = 100;      // service.cpp:56
  size_t min_read_length = 73;       // service.cpp:57
  std::string prefix, suffix, hashfile, pull, push, push_count, rocksdb_path, rocksdb_ext;
  std::vector<std::string> rocksdb;

  static WorkerConfig from(const Config& cfg) {
    WorkerConfig w;
    if (cfg.exists("size_of_sample")) w.size_of_sample = atoi(cfg.lookup_string("size_of_sample").c_str());
    if (cfg.exists("has_other_meta_data")) w.has_other_meta_data = cfg.lookup_string("has_other_meta_data") == "1";
    if (cfg.exists("max_read_length")) w.max_read_length = (size_t)atoi(cfg.lookup_string("max_read_length").c_str());
    if (cfg.exists("min_read_length")) w.min_read_length = (size_t)atoi(cfg.lookup_string("min_read_length").c_str());
    w.prefix = cfg.lookup_string("prefix");
    w.suffix = cfg.lookup_string("suffix");
    w.hashfile = cfg.lookup_string("hashfile");
    w.pull = cfg.lookup_string("pull");
    w.push = cfg.lookup_string("push");
    w.push_count = cfg.lookup_string("push_count");
    w.rocksdb_path = cfg.lookup_string("rocksdb_path");
    w.rocksdb_ext = cfg.lookup_string("rocksdb_ext");
    w.rocksdb = cfg.lookup_array("rocksdb");
    return w;
  }
  static WorkerConfig read_file(const std::string& path) { return from(Config::read_file(path)); }
};

// `name<TAB>code` lines of the hash file -> code -> name (service.cpp:1475-1488)
inline std::map<std::string, std::string> read_hash_file(const std::string& path) {
  std::map<std::string, std::string> code_to_name;
  std::ifstream f(path.c_str(), std::ios::binary);
  for (std::string l; std::getline(f, l);) {
    if (l.size() < 1) continue;
    const size_t tab = l.find('\t');
    if (tab == std::string::npos) continue;
    code_to_name.insert(std::make_pair(l.substr(tab + 1), l.substr(0, tab)));  // insert: the first line of a code wins
  }
  return code_to_name;
}

}  // namespace rsb200
#endif
