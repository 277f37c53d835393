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

}  // namespace rsb200

#endif
