// TEST PROGRAM for include/rs_wire.hpp (no GPU, no library): reads one command per line on stdin,
// tab-separated, and prints the encoded message as hex (or the decoded fields).
//   req    <hex>                                  -> "ok t rt q k s p a isalt" with '-' for absent, or "bad"
//   count  rt q revcomp count                     -> hex of Reply
//   reads  rt q revcomp read1,read2,...           -> hex
//   all    rt q revcomp read:g/c/l;g/c/l|read:... -> hex
//   samples rt q g/c/l;g/c/l                      -> hex
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>

#include "rs_wire.hpp"

using namespace rsb200::wire;

static std::string hex(const std::string& s) {
  static const char* d = "0123456789abcdef";
  std::string o;
  for (unsigned char c : s) { o.push_back(d[c >> 4]); o.push_back(d[c & 15]); }
  return o;
}
static std::string unhex(const std::string& h) {
  std::string o;
  for (size_t i = 0; i + 1 < h.size(); i += 2) o.push_back((char)strtol(h.substr(i, 2).c_str(), nullptr, 16));
  return o;
}
static std::vector<std::string> split(const std::string& s, char sep) {
  std::vector<std::string> v;
  if (s.empty()) return v;
  std::stringstream ss(s);
  std::string item;
  while (std::getline(ss, item, sep)) v.push_back(item);
  return v;
}
static std::vector<ReadInfo> infos(const std::string& s) {
  std::vector<ReadInfo> v;
  for (const std::string& one : split(s, ';')) {
    const std::vector<std::string> f = split(one, '/');
    ReadInfo i;
    i.g = f[0];
    i.c = atoi(f[1].c_str());
    i.l = atoi(f[2].c_str());
    v.push_back(i);
  }
  return v;
}

int main() {
  std::string line;
  while (std::getline(std::cin, line)) {
    std::vector<std::string> f = split(line, '\t');
    while (f.size() < 5) f.push_back("");
    if (f[0] == "req") {
      Request r;
      const std::string b = unhex(f[1]);
      if (!decode_request(b.data(), b.size(), r)) { std::cout << "bad\n"; continue; }
      std::cout << "ok " << r.t << " " << r.rt << " " << r.q << " " << (r.has_k ? std::to_string(r.k) : "-") << " "
                << (r.has_s ? std::to_string(r.s) : "-") << " " << (r.has_p ? std::to_string(r.p) : "-") << " "
                << (r.has_a ? r.a : "-") << " " << (r.has_isalt ? std::to_string(r.isalt) : "-") << " "
                << hex(encode_request(r)) << "\n";
    } else if (f[0] == "count") {
      std::cout << hex(encode_reply_count(atoi(f[1].c_str()), f[2], f[3] == "1", strtoull(f[4].c_str(), nullptr, 10))) << "\n";
    } else if (f[0] == "reads") {
      std::cout << hex(encode_reply_reads(atoi(f[1].c_str()), f[2], f[3] == "1", split(f[4], ','))) << "\n";
    } else if (f[0] == "all") {
      std::vector<ResultAll> hits;
      for (const std::string& h : split(f[4], '|')) {
        const size_t c = h.find(':');
        ResultAll a;
        a.r = h.substr(0, c);
        a.s = infos(h.substr(c + 1));
        hits.push_back(a);
      }
      std::cout << hex(encode_reply_all(atoi(f[1].c_str()), f[2], f[3] == "1", hits)) << "\n";
    } else if (f[0] == "samples") {
      std::cout << hex(encode_reply_samples(atoi(f[1].c_str()), f[2], infos(f[3]))) << "\n";
    }
  }
  return 0;
}
