// TEST PROGRAM: the reference's UNMODIFIED src/bwt/query.cpp (compiled from /root/reference by oracle/Makefile) linked
// against rsb200::GpuBWT -- the drop-in mode of INTEGRATION.md section 2: code that holds a `const BWT*` (every handler
// of the query worker) keeps running, each virtual getter a batch of one through the C ABI.
//   ref_query_over_gpubwt <file.bwt> <queries.txt> <rows.txt>
// Per query line:  "I <lower> <upper>" from the reference's findInterval, "X <0|1>" from its query_exactmatch and
// "Q <n> <sha-less listing>" = the reads its query() rebuilds (joined by ','); per row line: "R <prefix>|<postfix>".
#include <fstream>
#include <iostream>

#include "bwt.h"    // the reference's own headers (include/bwt), found through -I$(REF)/include/bwt
#include "query.h"

#include "rs_gpu_bwt.hpp"

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  try {
    rsb200::GpuBWT gpu(argv[1], 0);
    const BWT* p = &gpu;  // all the reference's query.cpp ever sees
    std::ifstream qf(argv[2]);
    for (std::string w; std::getline(qf, w);) {
      if (w.empty()) continue;
      const BWTInterval iv = findInterval(p, w);
      std::cout << "I " << iv.lower << " " << iv.upper << "\n";
      std::cout << "X " << (query_exactmatch(p, w) ? 1 : 0) << "\n";
      const std::vector<std::string> reads = query(p, w);
      std::cout << "Q " << reads.size();
      for (const std::string& r : reads) std::cout << " " << r;
      std::cout << "\n";
    }
    std::ifstream rf(argv[3]);
    for (size_t row; rf >> row;) std::cout << "R " << extractPrefix(p, row) << "|" << extractPostfix(p, row) << "\n";
  } catch (const std::exception& e) {
    std::cerr << "error: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
