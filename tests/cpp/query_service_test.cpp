// TEST PROGRAM for rsb200::QueryBatcher (include/rs_service.hpp):
//   query_service_test <file.bwt> <requests.hex> <merged.rlosorted.reads> <hashfile> <size_of_sample>
// Every line of the second file is one serialised Request in hex; the third is the ETL's merged read file the
// sample table is loaded from, the fourth the `name<TAB>code` hash file (service.cpp:1475-1488).
// Prints "skip" for a message the batcher does not take, then one hex line per Reply.
#include <cstdlib>
#include <fstream>
#include <iostream>

#include "rs_service.hpp"

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

int main(int argc, char** argv) {
  if (argc < 6) return 2;
  try {
    const int sos = atoi(argv[5]);
    rsb200::GpuBWT bwt(argv[1], 0);
    bwt.loadSamplesFile(argv[3], sos, false);
    std::map<std::string, std::string> code_to_name;
    std::ifstream hf(argv[4]);
    for (std::string l; std::getline(hf, l);) {
      const size_t tab = l.find('\t');
      if (tab != std::string::npos) code_to_name[l.substr(tab + 1)] = l.substr(0, tab);
    }
    rsb200::QueryBatcher batcher(&bwt, &code_to_name, sos);
    std::ifstream in(argv[2]);
    for (std::string line; std::getline(in, line);)
      if (!batcher.push(unhex(line))) std::cout << "skip\n";
    std::cout << "pending " << batcher.pending() << "\n";
    const size_t n = batcher.flush([](const std::string& bytes) { std::cout << hex(bytes) << "\n"; });
    std::cout << "sent " << n << "\n";
  } catch (const std::exception& e) {
    std::cerr << "error: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
