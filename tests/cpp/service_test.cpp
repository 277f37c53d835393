// TEST PROGRAM for include/rs_service.hpp: service_test <file.bwt> <requests.hex>
// Every line of the second file is one serialised Request in hex (what the SUB socket would deliver).
// Prints "skip" for a message the count batcher does not take, then after the batch one hex line per Reply.
#include <cstdio>
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
  if (argc < 3) return 2;
  try {
    rsb200::GpuBWT bwt(argv[1], 0);
    rsb200::CountBatcher batcher(&bwt);
    std::ifstream in(argv[2]);
    std::string line;
    while (std::getline(in, line))
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
