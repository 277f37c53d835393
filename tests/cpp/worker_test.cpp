// TEST PROGRAM for rsb200::Worker (include/rs_service.hpp): the receive loop over an in-memory transport.
//   worker_test <worker.cfg> <requests.hex> <merged.rlosorted.reads> [<sai: raw uint64, one per read>]
// With the fourth argument SiteMatch requests are answered too (SiteBatcher; the aligner here is a stand-in predicate,
// the reference's align_gt_read being host code that stays the reference's).
// The .cfg names the .bwt (prefix), the hash file, size_of_sample, has_other_meta_data, the read lengths and the shard's
// suffix, as the reference's worker reads them (service.cpp:1386-1442); every line of the second file is one serialised
// Request in hex.  Prints one line per outgoing message: "push <hex>", "count <hex>" or "other <hex of the request>".
#include <cstdlib>
#include <deque>
#include <fstream>
#include <iostream>

#include "rs_config.hpp"
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

struct QueueTransport {
  std::deque<std::string> in;
  bool recv(std::string& bytes, std::chrono::microseconds) {
    if (in.empty()) return false;
    bytes.swap(in.front());
    in.pop_front();
    return true;
  }
  void send(const std::string& b) { std::cout << "push " << hex(b) << "\n"; }
  void send_count(const std::string& b) { std::cout << "count " << hex(b) << "\n"; }
  bool stopped() { return in.empty(); }
};

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  try {
    const rsb200::WorkerConfig cfg = rsb200::WorkerConfig::read_file(argv[1]);
    rsb200::GpuBWT bwt(cfg.prefix + ".bwt", 0);
    bwt.loadSamplesFile(argv[3], cfg.size_of_sample, cfg.has_other_meta_data);
    const std::map<std::string, std::string> code_to_name = rsb200::read_hash_file(cfg.hashfile);
    rsb200::CountBatcher counts(&bwt, 64, std::chrono::microseconds(1000000));
    rsb200::QueryBatcher queries(&bwt, &code_to_name, cfg.size_of_sample, cfg.min_read_length, 16, 256,
                                 cfg.has_other_meta_data, cfg.suffix, cfg.max_read_length, std::chrono::microseconds(1000000));
    const bool sites_on = argc > 4;
    if (sites_on) {
      std::ifstream sf(argv[4], std::ios::binary);
      std::vector<uint64_t> sai;
      for (uint64_t v; sf.read(reinterpret_cast<char*>(&v), 8);) sai.push_back(v);
      bwt.setSai(sai);
      bwt.loadReadsFile(argv[3]);
    }
    auto align = [](const std::string& read, const std::string& window, size_t pos, const std::string& alt, bool isalt) {
      return (read.size() + (unsigned char)read[read.size() / 2] + window.size() + pos + alt.size() + (isalt ? 1 : 0)) % 3 != 0;
    };
    rsb200::SiteBatcher<decltype(align)> sites(&bwt, &code_to_name, cfg.size_of_sample, align, cfg.max_read_length, 64, 256,
                                               cfg.has_other_meta_data);
    QueueTransport io;
    std::ifstream in(argv[2]);
    for (std::string line; std::getline(in, line);) io.in.push_back(unhex(line));
    auto other = [&](const std::string& bytes) {
      if (!(sites_on && sites.push(bytes))) std::cout << "other " << hex(bytes) << "\n";
    };
    rsb200::Worker<QueueTransport, decltype(other)> worker(&io, &counts, &queries, other);
    worker.run();
    const size_t site_replies = sites.flush([&io](const std::string& b) { io.send(b); });
    std::cout << "sent " << worker.replies_sent() + site_replies << "\n";
  } catch (const std::exception& e) {
    std::cerr << "error: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
