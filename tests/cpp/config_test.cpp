// TEST PROGRAM for include/rs_config.hpp and rsb200::BatchWindow (include/rs_service.hpp); no GPU, no library.
//   config_test parse <file.cfg> key...        prints  key<TAB>S<TAB>value | key<TAB>A<TAB>v1|v2|.. | key<TAB>-  per key
//   config_test worker <file.cfg>              prints the WorkerConfig fields
//   config_test window                         prints the BatchWindow decisions of a scripted timeline
#include <iostream>

#include "rs_config.hpp"
#include "rs_service.hpp"

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const std::string mode = argv[1];
  try {
    if (mode == "parse") {
      const rsb200::Config c = rsb200::Config::read_file(argv[2]);
      for (int i = 3; i < argc; ++i) {
        const std::string k = argv[i];
        if (!c.exists(k)) {
          std::cout << k << "\t-\n";
          continue;
        }
        try {
          std::string v = c.lookup_string(k);
          for (char& ch : v)
            if (ch == '\t') ch = '\x01';  // keep the line format
          std::cout << k << "\tS\t" << v << "\n";
        } catch (const rsb200::ConfigError&) {
          try {
            const std::vector<std::string>& a = c.lookup_array(k);
            std::cout << k << "\tA\t";
            for (size_t j = 0; j < a.size(); ++j) std::cout << (j ? "|" : "") << a[j];
            std::cout << "\n";
          } catch (const rsb200::ConfigError&) {
            std::cout << k << "\tX\n";
          }
        }
      }
    } else if (mode == "worker") {
      const rsb200::WorkerConfig w = rsb200::WorkerConfig::read_file(argv[2]);
      std::cout << w.size_of_sample << " " << w.has_other_meta_data << " " << w.max_read_length << " " << w.min_read_length
                << " " << w.prefix << " [" << w.suffix << "] " << w.hashfile << " " << w.pull << " " << w.push << " "
                << w.push_count << " " << w.rocksdb_path << " " << w.rocksdb_ext << " " << w.rocksdb.size() << "\n";
    } else if (mode == "window") {
      typedef rsb200::BatchWindow::clock clock;
      const clock::time_point t0 = clock::now();
      auto at = [&](int us) { return t0 + std::chrono::microseconds(us); };
      rsb200::BatchWindow w(3, std::chrono::microseconds(1000));
      std::cout << w.due(0, at(0)) << w.wait_budget(0, at(0)).count() << " ";  // nothing pending: not due, full wait
      w.on_push(1, at(100));
      std::cout << w.due(1, at(500)) << " " << w.wait_budget(1, at(500)).count() << " ";  // 400 us in: 600 us left
      std::cout << w.due(1, at(1100)) << " " << w.wait_budget(1, at(1200)).count() << " ";  // window passed
      w.on_push(2, at(1150));
      w.on_push(3, at(1160));
      std::cout << w.due(3, at(1160)) << " ";  // size limit
      w.on_flush();
      std::cout << w.due(0, at(5000)) << " ";
      w.on_push(1, at(6000));  // the window restarts with the next batch's first request
      std::cout << w.due(1, at(6500)) << w.due(1, at(7000)) << "\n";
    }
  } catch (const rsb200::ConfigError& e) {
    std::cout << "error: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
