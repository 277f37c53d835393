// TEST PROGRAM for include/rs_gpu_bwt.hpp: drives the C++ mirror of the reference interface the way
// the reference's worker does -- through a `const BWT*` (scalar virtual calls) and through the batched
// replacements -- and prints the results for tests/test_cpp_shim.py to compare with the oracle.
//   usage: shim_test <file.bwt> <queries.txt> [file.sai.u64]
#include <fstream>
#include <iostream>

#include "rs_gpu_bwt.hpp"

// findInterval exactly as the reference's callers reach it: virtual calls on the abstract class
// (updateInterval / initInterval, include/bwt/query.h:18-24).
static BWTInterval scalar_find_interval(const BWT* p, const std::string& w) {
  int j = (int)w.size() - 1;
  BWTInterval iv;
  iv.lower = p->getPC(w[j]);
  iv.upper = iv.lower + p->getOcc(w[j], p->getBWLen() - 1) - 1;
  for (--j; j >= 0; --j) {
    const uint64_t pb = p->getPC(w[j]);
    iv.lower = pb + p->getOcc(w[j], iv.lower - 1);
    iv.upper = pb + p->getOcc(w[j], iv.upper) - 1;
    if (iv.lower > iv.upper) break;
  }
  return iv;
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  try {
    rsb200::GpuBWT bwt(argv[1]);
    const BWT* p = &bwt;
    std::vector<std::string> qs;
    std::ifstream in(argv[2]);
    for (std::string l; std::getline(in, l);)
      if (!l.empty()) qs.push_back(l);
    std::cout << "bwlen " << p->getBWLen() << " pc " << p->getPC('$') << ' ' << p->getPC('A') << ' ' << p->getPC('C') << ' '
              << p->getPC('G') << ' ' << p->getPC('T') << "\n";
    const std::vector<BWTInterval> ivs = rsb200::findIntervalBatch(&bwt, qs);
    for (size_t i = 0; i < qs.size(); ++i) {
      const BWTInterval s = i < 8 ? scalar_find_interval(p, qs[i]) : ivs[i];
      const std::vector<std::string> reads = i < 16 ? rsb200::query(&bwt, qs[i], 256) : std::vector<std::string>();
      std::cout << "q " << ivs[i].lower << ' ' << ivs[i].upper << ' ' << s.lower << ' ' << s.upper << ' ' << reads.size() << ' '
                << (i < 16 ? (int)rsb200::query_exactmatch(&bwt, qs[i], 256) : -1);
      for (const std::string& r : reads) std::cout << ' ' << r;
      std::cout << "\n";
    }
    std::cout << "f " << p->getF(0) << p->getF(p->getPC('C')) << p->getF(p->getBWLen() - 1) << " char " << p->getChar(12345)
              << " occat " << p->getOccAt('G', 777) << "\n";
    const std::vector<std::vector<BWTInterval> > tiles = rsb200::tileIntervalsBatch(&bwt, qs, 31, 4);
    for (size_t i = 0; i < qs.size(); ++i) {
      std::cout << "tiles " << tiles[i].size();
      for (const BWTInterval& iv : tiles[i]) std::cout << ' ' << iv.lower << ' ' << iv.upper;
      std::cout << "\n";
    }
    if (argc > 3) {  // /gt seed stage: every query is a window with the site in its middle
      std::ifstream sf(argv[3], std::ios::binary);
      std::vector<uint64_t> sai;
      for (uint64_t v; sf.read(reinterpret_cast<char*>(&v), 8);) sai.push_back(v);
      bwt.setSai(sai);
      std::vector<uint64_t> pos;
      for (size_t i = 0; i < qs.size(); ++i) pos.push_back(qs[i].size() / 2 + 1);
      std::vector<uint32_t> flags;
      const std::vector<std::vector<uint64_t> > gt = rsb200::findGtReadsBatch(&bwt, qs, pos, 31, 7, 12, &flags);
      for (size_t i = 0; i < qs.size(); ++i) {
        std::cout << "gt " << flags[i] << ' ' << gt[i].size();
        for (size_t j = 0; j < gt[i].size(); ++j) std::cout << ' ' << gt[i][j];
        std::cout << "\n";
      }
    }
    try {
      rsb200::GpuBWT bad("/nonexistent.bwt");
      std::cout << "error none\n";
    } catch (const rsb200::Error& e) {
      std::cout << "error " << e.status << "\n";
    }
  } catch (const std::exception& e) {
    std::cerr << "FAILED: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
