// rs_gpu_bwt.hpp -- C++ host-side mirror of the reference's BWT interface over the C ABI.
//
// `GpuBWT` derives from the reference's abstract `class BWT` (/root/reference/include/bwt/bwt.h:6-15)
// so that code holding a `const BWT*` -- every task of the query worker, src/service/service.cpp:1449,
// 1513 -- keeps compiling and linking: each virtual getter is a batch of one through the C ABI
// (include/readserver_b200.h).  That path is for the handlers that stay scalar (the /gt interval
// extension loop, service.cpp:523-705); the hot handlers call the *Batch functions below instead
// (count_reads service.cpp:279-315, find_reads :714-797), which is where the GPU pays off.
//
// Build: include the reference's own "bwt.h" / "query.h" BEFORE this header inside the reference tree
// (they define RS_BWT_H / RS_BWT_QUERY_H).  Outside it, equivalent declarations are provided here so
// the header is self-contained.  Link with -lrs_b200 (readserver_b200/librs_b200.so).
#ifndef RS_GPU_BWT_HPP
#define RS_GPU_BWT_HPP

#include <stdint.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "readserver_b200.h"

#ifndef RS_BWT_H
#define RS_BWT_H
// Same abstract interface as the reference's include/bwt/bwt.h (used only when that header is absent).
class BWT {
 public:
  virtual char getChar(const uint64_t& index) const = 0;
  virtual uint64_t getOccAt(const char& b, const uint64_t& index) const = 0;
  virtual uint64_t getOcc(const char& b, const uint64_t& index) const = 0;
  virtual char getF(const uint64_t& index) const = 0;
  virtual uint64_t getPC(const char& b) const = 0;
  virtual uint64_t getBWLen() const = 0;
};
#endif

#ifndef RS_BWT_QUERY_H
#define RS_BWT_QUERY_H
struct BWTInterval {  // include/bwt/query.h:8-11
  uint64_t lower;
  uint64_t upper;
};
#endif

namespace rsb200 {

struct Error : std::runtime_error {
  int status;
  Error(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

inline void check(int rc) {
  if (rc != RS_STATUS_OK) throw Error(rc, rs_last_error());
}

// One BWT shard resident in the HBM of one GPU.  Immutable after construction (+ optional tables);
// safe to share between threads like the reference's `const BWT*` (service.cpp:88-89, 1505-1513).
class GpuBWT : public BWT {
 public:
  // RLEBWT(filename), src/bwt/rlebwt.cpp:11-32.  A bad magic number makes the reference exit(1)
  // (rlebwt_reader.cpp:31-34); here every failure throws rsb200::Error instead.
  explicit GpuBWT(const std::string& bwt_path, int device = 0) : ix_(nullptr) {
    check(rs_index_open(bwt_path.c_str(), device, &ix_));
  }
  GpuBWT(const uint8_t* runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols, int device = 0)
      : ix_(nullptr) {
    check(rs_index_from_runs(runs, num_runs, num_strings, num_symbols, device, &ix_));
  }
  ~GpuBWT() { rs_index_close(ix_); }
  GpuBWT(const GpuBWT&) = delete;
  GpuBWT& operator=(const GpuBWT&) = delete;

  rs_index* handle() const { return ix_; }
  void setSai(const std::vector<uint64_t>& sai) { check(rs_index_set_sai(ix_, sai.data(), sai.size())); }
  void setSamples(const std::vector<uint64_t>& offsets, const std::vector<uint8_t>& codes, int size_of_sample) {
    check(rs_index_set_samples(ix_, offsets.data(), codes.data(), offsets.size() - 1, size_of_sample));
  }

  // the same table from the ETL's merged `read / codes` file (load_data_into_rocksdb.cpp:41-49)
  void loadSamplesFile(const std::string& path, int size_of_sample, bool has_other_meta_data) {
    check(rs_index_load_samples_file(ix_, path.c_str(), size_of_sample, has_other_meta_data ? 1 : 0));
  }
  // optional locate table (8 bytes of HBM per BWT symbol): the locate stage reads one entry per hit row instead of
  // walking; results do not depend on it
  // every > 1: sampled variant, n (1/8 + 1/64 + 8/every) bytes, a hit row is walked fewer than `every` steps
  void buildLocateTable(uint32_t max_read_len = 4096, uint32_t every = 1) {
    check(rs_index_build_locate_samples(ix_, max_read_len, every));
  }
  void dropLocateTable() { check(rs_index_drop_locate_table(ix_)); }
  // packed read table for output=reads, from the same file (reads holding 'N' are not in the BWT)
  void loadReadsFile(const std::string& path) { check(rs_index_load_reads_file(ix_, path.c_str())); }

  // ---- class BWT: batches of one ----
  char getChar(const uint64_t& index) const override {
    char c;
    check(rs_get_char_batch(ix_, &index, 1, &c));
    return c;
  }
  uint64_t getOccAt(const char& b, const uint64_t& count) const override {
    uint64_t r;
    check(rs_get_occ_at_batch(ix_, &b, &count, 1, &r));
    return r;
  }
  uint64_t getOcc(const char& b, const uint64_t& index) const override {
    uint64_t r;
    check(rs_get_occ_batch(ix_, &b, &index, 1, &r));
    return r;
  }
  char getF(const uint64_t& index) const override {
    char c;
    check(rs_get_f_batch(ix_, &index, 1, &c));
    return c;
  }
  uint64_t getPC(const char& b) const override { return rs_get_pc(ix_, b); }
  uint64_t getBWLen() const override { return rs_get_bwlen(ix_); }

 private:
  rs_index* ix_;
};

// ---- include/bwt/query.h:18-32, batched -------------------------------------------------------------
namespace detail {
inline void flatten(const std::vector<std::string>& ws, std::string& text, std::vector<uint64_t>& offs) {
  offs.assign(ws.size() + 1, 0);
  size_t total = 0;
  for (size_t i = 0; i < ws.size(); ++i) {
    total += ws[i].size();
    offs[i + 1] = total;
  }
  text.clear();
  text.reserve(total + 1);
  for (const std::string& w : ws) text += w;
}
}  // namespace detail

// findInterval (query.cpp:24-41) for many strings of any length.
inline std::vector<BWTInterval> findIntervalBatch(const GpuBWT* bwt, const std::vector<std::string>& ws) {
  std::string text;
  std::vector<uint64_t> offs;
  detail::flatten(ws, text, offs);
  std::vector<uint64_t> lo(ws.size()), up(ws.size());
  if (!ws.empty()) check(rs_find_interval_var_batch(bwt->handle(), text.data(), offs.data(), ws.size(), lo.data(), up.data()));
  std::vector<BWTInterval> out(ws.size());
  for (size_t i = 0; i < ws.size(); ++i) {
    out[i].lower = lo[i];
    out[i].upper = up[i];
  }
  return out;
}

inline BWTInterval findInterval(const GpuBWT* bwt, const std::string& w) {
  return findIntervalBatch(bwt, std::vector<std::string>(1, w))[0];
}

// count_reads (service.cpp:279-315) for many k-mers of one length; strand 0 forward, 1 reverse complement.
inline std::vector<uint64_t> countReadsBatch(const GpuBWT* bwt, const std::string& kmers, int k, int strand) {
  std::vector<uint64_t> c(kmers.size() / (size_t)k);
  if (!c.empty()) check(rs_count_batch(bwt->handle(), kmers.data(), k, c.size(), strand, c.data()));
  return c;
}

// extractPrefix / extractPostfix (query.cpp:43-85) for many rows; max_len bounds the walk (a walk inside
// the reference's tail region can cycle: such rows come back as empty strings with ok[i] = false).
inline std::vector<std::string> extractPrefixBatch(const GpuBWT* bwt, const std::vector<uint64_t>& rows, uint32_t max_len,
                                                   std::vector<uint64_t>* lexrank = nullptr, std::vector<bool>* ok = nullptr) {
  const size_t n = rows.size();
  std::vector<uint64_t> lex(n);
  std::vector<uint32_t> st(n);
  std::vector<char> bases(n * (size_t)max_len);
  if (n) check(rs_extract_prefix_batch(bwt->handle(), rows.data(), n, max_len, lex.data(), st.data(), nullptr, bases.data()));
  std::vector<std::string> out(n);
  if (ok) ok->assign(n, true);
  for (size_t i = 0; i < n; ++i) {
    if (st[i] == 0xFFFFFFFFu) {
      if (ok) (*ok)[i] = false;
    } else {
      out[i].assign(bases.data() + i * (size_t)max_len, st[i]);
    }
  }
  if (lexrank) lexrank->swap(lex);
  return out;
}

inline std::vector<std::string> extractPostfixBatch(const GpuBWT* bwt, const std::vector<uint64_t>& rows, uint32_t max_len,
                                                    std::vector<uint64_t>* read_index = nullptr,
                                                    std::vector<bool>* ok = nullptr) {
  const size_t n = rows.size();
  std::vector<uint64_t> ri(n);
  std::vector<uint32_t> st(n);
  std::vector<char> bases(n * (size_t)max_len);
  if (n) check(rs_extract_postfix_batch(bwt->handle(), rows.data(), n, max_len, ri.data(), st.data(), bases.data()));
  std::vector<std::string> out(n);
  if (ok) ok->assign(n, true);
  for (size_t i = 0; i < n; ++i) {
    if (st[i] == 0xFFFFFFFFu) {
      if (ok) (*ok)[i] = false;
    } else {
      out[i].assign(bases.data() + i * (size_t)max_len, st[i]);
    }
  }
  if (read_index) read_index->swap(ri);
  return out;
}

inline std::string extractPrefix(const GpuBWT* bwt, const size_t& index, uint32_t max_len = 1024) {
  return extractPrefixBatch(bwt, std::vector<uint64_t>(1, index), max_len)[0];
}
inline std::string extractPostfix(const GpuBWT* bwt, const size_t& index, uint32_t max_len = 1024) {
  return extractPostfixBatch(bwt, std::vector<uint64_t>(1, index), max_len)[0];
}

// query (query.cpp:87-100): every read containing w, rebuilt as extractPrefix + extractPostfix per row of
// w's interval; a string with a byte outside ACGT yields nothing (query.cpp:90).
inline std::vector<std::string> query(const GpuBWT* bwt, const std::string& w, uint32_t max_len = 1024) {
  std::vector<std::string> out;
  if (w.empty() || w.find_first_not_of("ACGT") != std::string::npos) return out;
  const BWTInterval iv = findInterval(bwt, w);
  if (iv.upper < iv.lower) return out;
  std::vector<uint64_t> rows;
  for (uint64_t i = iv.lower; i <= iv.upper; ++i) rows.push_back(i);
  const std::vector<std::string> pre = extractPrefixBatch(bwt, rows, max_len);
  const std::vector<std::string> post = extractPostfixBatch(bwt, rows, max_len);
  for (size_t i = 0; i < rows.size(); ++i) out.push_back(pre[i] + post[i]);
  return out;
}

// The reads behind read indices (as the locate stage returns them), from the packed read table: what
// query() rebuilds per row with extractPrefix + extractPostfix.  An index outside the table gives "".
inline std::vector<std::string> readsBatch(const GpuBWT* bwt, const std::vector<uint64_t>& read_index, uint32_t max_len = 256) {
  std::vector<std::string> out(read_index.size());
  if (read_index.empty()) return out;
  std::vector<char> buf(read_index.size() * (size_t)max_len);
  std::vector<uint32_t> len(read_index.size());
  check(rs_reads_batch(bwt->handle(), read_index.data(), read_index.size(), max_len, buf.data(), len.data()));
  for (size_t i = 0; i < out.size(); ++i)
    if (len[i] != 0xFFFFFFFFu) out[i].assign(&buf[i * (size_t)max_len], len[i] < max_len ? len[i] : max_len);
  return out;
}

// get_tiles + findInterval per tile (service.cpp:232-246, 466-500) for a batch of queries: tiles of `kmer`
// bases every skip+1 bases; result[q] = the intervals of query q's tiles, (1, 0) for a tile the reference skips.
inline std::vector<std::vector<BWTInterval> > tileIntervalsBatch(const GpuBWT* bwt, const std::vector<std::string>& ws,
                                                                size_t kmer, size_t skip) {
  std::string text;
  std::vector<uint64_t> offs;
  detail::flatten(ws, text, offs);
  uint64_t cap = 0;
  for (const std::string& w : ws) cap += w.size() < kmer ? 0 : (w.size() - kmer) / (skip + 1) + 1;
  std::vector<uint64_t> toff(ws.size() + 1), lo(cap ? cap : 1), up(cap ? cap : 1);
  check(rs_tile_intervals_batch(bwt->handle(), text.data(), offs.data(), ws.size(), (uint32_t)kmer, (uint32_t)skip,
                                toff.data(), cap, lo.data(), up.data()));
  std::vector<std::vector<BWTInterval> > out(ws.size());
  for (size_t q = 0; q < ws.size(); ++q)
    for (uint64_t t = toff[q]; t < toff[q + 1]; ++t) {
      BWTInterval iv;
      iv.lower = lo[t];
      iv.upper = up[t];
      out[q].push_back(iv);
    }
  return out;
}

// find_gt_reads over every tile of GtTask::run's loop (service.cpp:507-711, 1044-1072) for a batch of site
// queries: ws[q] is the oriented window cut around the site, pos[q] the site's 1-based position in it.
// result[q] = the sorted, unique read indices (builder input order) of the reads the reference would hand to the
// aligner; flags[q] (optional) receives RS_GT_FLAG_* for windows on which the reference itself throws or hangs.
inline std::vector<std::vector<uint64_t> > findGtReadsBatch(const GpuBWT* bwt, const std::vector<std::string>& ws,
                                                            const std::vector<uint64_t>& pos, size_t kmer, size_t skip,
                                                            uint64_t max_interval_size = 10000,
                                                            std::vector<uint32_t>* flags = 0, uint32_t max_steps = 1024,
                                                            uint64_t max_rows = (uint64_t)1 << 26) {
  std::string text;
  std::vector<uint64_t> offs;
  detail::flatten(ws, text, offs);
  std::vector<uint64_t> ooff(ws.size() + 1, 0), ri((size_t)1 << 12);
  std::vector<uint32_t> fl(ws.size() ? ws.size() : 1, 0);
  uint64_t cand = 0;
  for (;;) {
    const int rc = rs_gt_reads_batch(bwt->handle(), text.data(), offs.data(), pos.data(), ws.size(), (uint32_t)kmer,
                                     (uint32_t)skip, max_interval_size, max_steps, max_rows, ooff.data(), ri.size(),
                                     ri.data(), fl.data(), &cand);
    if (rc == RS_STATUS_CAPACITY && cand <= max_rows && ooff[ws.size()] > ri.size()) {
      ri.resize(ooff[ws.size()]);
      continue;
    }
    check(rc);
    break;
  }
  std::vector<std::vector<uint64_t> > out(ws.size());
  for (size_t q = 0; q < ws.size(); ++q) out[q].assign(ri.begin() + ooff[q], ri.begin() + ooff[q + 1]);
  if (flags) flags->assign(fl.begin(), fl.begin() + ws.size());
  return out;
}

// query_exactmatch (query.cpp:102-120): is w one of the reads?
inline bool query_exactmatch(const GpuBWT* bwt, const std::string& w, uint32_t max_len = 1024) {
  for (const std::string& s : query(bwt, w, max_len))
    if (s == w) return true;
  return false;
}

}  // namespace rsb200
#endif
