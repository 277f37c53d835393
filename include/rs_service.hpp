// rs_service.hpp -- the count path of the reference's query worker, batched.
//
// The reference worker handles one message at a time: `Request` -> count_reads(Forward) and
// count_reads(RevComp), each building and sending one `Reply` (service.cpp:279-315, 1548-1575).  A GPU
// index wants thousands of queries per launch, so this shell collects the count requests of one receive
// window, searches them with rs_count_batch (one call per query length and strand) and emits exactly the
// replies the reference would have sent, in arrival order, forward before reverse complement.
//
// Transport stays outside: `push()` takes the bytes a ZMQ SUB socket delivered, `flush()` hands each
// serialised Reply to a callable (the PUSH socket's send).  libzmq is not in this image, so the socket loop
// itself is the maintainer's three lines around these two calls.
#ifndef RS_SERVICE_HPP
#define RS_SERVICE_HPP

#include <map>
#include <string>
#include <utility>
#include <vector>

#include "rs_gpu_bwt.hpp"
#include "rs_wire.hpp"

namespace rsb200 {

class CountBatcher {
 public:
  explicit CountBatcher(const GpuBWT* bwt, size_t max_pending = (size_t)1 << 20) : bwt_(bwt), max_pending_(max_pending) {}

  // One message from the SUB socket.  Returns true when it was a count request (CountReads, or ExactMatch
  // with output=count: the two cases the reference answers with count_reads) and has been queued; false
  // when the caller must route it elsewhere (other request types, or bytes that do not parse).
  bool push(const void* data, size_t size) {
    wire::Request r;
    if (!wire::decode_request(data, size, r)) return false;
    int32_t rt;
    if (r.t == wire::CountReads) rt = wire::CountReads;
    else if (r.t == wire::ExactMatch && r.rt == wire::Count) rt = wire::ExactMatch;
    else return false;
    pending_.push_back(Item{std::move(r.q), rt});
    return true;
  }
  bool push(const std::string& bytes) { return push(bytes.data(), bytes.size()); }

  size_t pending() const { return pending_.size(); }
  bool full() const { return pending_.size() >= max_pending_; }

  // Searches everything queued and calls send(const std::string&) once per Reply.  Returns the number of
  // replies (two per request).
  template <class Send>
  size_t flush(Send&& send) {
    const size_t n = pending_.size();
    if (!n) return 0;
    std::vector<uint64_t> fwd(n, 0), rev(n, 0);
    // one batch per query length; an empty query (undefined in the reference: query.cpp:25-27) counts 0
    std::map<size_t, std::vector<size_t>> by_len;
    for (size_t i = 0; i < n; ++i)
      if (!pending_[i].q.empty()) by_len[pending_[i].q.size()].push_back(i);
    for (const auto& kv : by_len) {
      std::string kmers;
      kmers.reserve(kv.first * kv.second.size());
      for (size_t i : kv.second) kmers += pending_[i].q;
      const std::vector<uint64_t> f = countReadsBatch(bwt_, kmers, (int)kv.first, 0);
      const std::vector<uint64_t> r = countReadsBatch(bwt_, kmers, (int)kv.first, 1);
      for (size_t j = 0; j < kv.second.size(); ++j) {
        fwd[kv.second[j]] = f[j];
        rev[kv.second[j]] = r[j];
      }
    }
    for (size_t i = 0; i < n; ++i) {
      send(wire::encode_reply_count(pending_[i].rt, pending_[i].q, false, fwd[i]));
      send(wire::encode_reply_count(pending_[i].rt, pending_[i].q, true, rev[i]));
    }
    pending_.clear();
    return 2 * n;
  }

 private:
  struct Item {
    std::string q;
    int32_t rt;
  };
  const GpuBWT* bwt_;
  size_t max_pending_;
  std::vector<Item> pending_;
};

// The ExactMatch branch with output = reads / all (QueryTask::run, service.cpp:1259-1362) for queries shorter than
// MIN_READ_LENGTH -- the k-mer lookups; longer queries go through the tile logic of find_reads (:755-794) and stay
// with the caller.  Per request and strand: findInterval of the query (as given: find_reads does not screen
// bytes), the rows in the order find_reads emits them (:724-749: the rows past the 2,048-row chunks first, then
// the chunks), each row's read as extractPrefix + extractPostfix, and for output = all the sample codes of the
// read (RocksDB value, :1325-1347) from the read-index-addressed sample table, names through the hash file's
// code -> name map.  One Reply per strand, forward first, bytes as QueryTask serialises them.
// Stores with has_other_meta_data are not covered: the sample table keeps codes only, c and l are sent as 0.
class QueryBatcher {
 public:
  QueryBatcher(const GpuBWT* bwt, const std::map<std::string, std::string>* code_to_name, int size_of_sample,
               size_t min_read_length = 73, size_t max_pending = (size_t)1 << 14, uint32_t max_len = 256)
      : bwt_(bwt), hash_(code_to_name), sos_(size_of_sample), min_len_(min_read_length), max_pending_(max_pending),
        max_len_(max_len) {}

  bool push(const void* data, size_t size) {
    wire::Request r;
    if (!wire::decode_request(data, size, r)) return false;
    if (r.t != wire::ExactMatch || r.rt == wire::Count || r.q.empty() || r.q.size() >= min_len_) return false;
    pending_.push_back(Item{std::move(r.q), r.rt});
    return true;
  }
  bool push(const std::string& bytes) { return push(bytes.data(), bytes.size()); }
  size_t pending() const { return pending_.size(); }
  bool full() const { return pending_.size() >= max_pending_; }

  static std::string rev_comp(const std::string& w) {  // service.cpp:253-276: other bytes are kept
    std::string s(w.rbegin(), w.rend());
    for (char& c : s) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'C' ? 'G' : c == 'G' ? 'C' : c;
    return s;
  }

  template <class Send>
  size_t flush(Send&& send) {
    const size_t n = pending_.size();
    if (!n) return 0;
    std::vector<std::string> ws;  // 2i: forward, 2i+1: reverse complement
    for (size_t i = 0; i < n; ++i) {
      ws.push_back(pending_[i].q);
      ws.push_back(rev_comp(pending_[i].q));
    }
    const std::vector<BWTInterval> ivs = findIntervalBatch(bwt_, ws);
    // rows of every search in find_reads' order
    const uint64_t large = 2048;  // large_match_size, service.cpp:86
    std::vector<uint64_t> rows, first(ws.size() + 1, 0);
    for (size_t j = 0; j < ws.size(); ++j) {
      const BWTInterval& iv = ivs[j];
      if (iv.upper >= iv.lower) {
        uint64_t start = iv.lower;
        while (iv.upper - start > 2 * large) start += large;
        for (uint64_t r = start; r <= iv.upper; ++r) rows.push_back(r);
        for (uint64_t r = iv.lower; r < start; ++r) rows.push_back(r);
      }
      first[j + 1] = rows.size();
    }
    std::vector<uint64_t> read_index;
    const std::vector<std::string> pre = extractPrefixBatch(bwt_, rows, max_len_);
    const std::vector<std::string> post = extractPostfixBatch(bwt_, rows, max_len_, &read_index);
    bool want_samples = false;
    for (size_t i = 0; i < n; ++i) want_samples |= pending_[i].rt != wire::Reads;
    std::vector<uint64_t> soff(rows.size() + 1, 0);
    std::vector<uint8_t> codes;
    if (want_samples && !rows.empty()) {
      codes.resize(1);
      int rc = rs_samples_batch(bwt_->handle(), read_index.data(), rows.size(), soff.data(), 0, codes.data());
      if (rc == RS_STATUS_CAPACITY) {
        codes.resize((size_t)soff[rows.size()] * sos_ + 1);
        rc = rs_samples_batch(bwt_->handle(), read_index.data(), rows.size(), soff.data(), soff[rows.size()], codes.data());
      }
      check(rc);
    }
    for (size_t i = 0; i < n; ++i)
      for (int rev = 0; rev < 2; ++rev) {
        const size_t j = 2 * i + rev;
        if (pending_[i].rt == wire::Reads) {
          std::vector<std::string> reads;
          for (uint64_t h = first[j]; h < first[j + 1]; ++h) reads.push_back(pre[h] + post[h]);
          send(wire::encode_reply_reads(wire::ExactMatch, pending_[i].q, rev != 0, reads));
        } else {
          std::vector<wire::ResultAll> hits;
          for (uint64_t h = first[j]; h < first[j + 1]; ++h) {
            wire::ResultAll one;
            one.r = pre[h] + post[h];
            for (uint64_t c = soff[h]; c < soff[h + 1]; ++c) {
              wire::ReadInfo info;
              const std::string code((const char*)&codes[(size_t)c * sos_], (size_t)sos_);
              const std::map<std::string, std::string>::const_iterator it = hash_->find(code);
              if (it != hash_->end()) info.g = it->second;  // (*hash)[code]: an unknown code reads as ""
              one.s.push_back(info);
            }
            hits.push_back(one);
          }
          send(wire::encode_reply_all(wire::ExactMatch, pending_[i].q, rev != 0, hits, pending_[i].rt));
        }
      }
    pending_.clear();
    return 2 * n;
  }

 private:
  struct Item {
    std::string q;
    int32_t rt;
  };
  const GpuBWT* bwt_;
  const std::map<std::string, std::string>* hash_;
  int sos_;
  size_t min_len_, max_pending_;
  uint32_t max_len_;
  std::vector<Item> pending_;
};

}  // namespace rsb200

#endif
