// rs_service.hpp -- the reference's query worker around the batched GPU calls.
//
// The reference worker handles one message at a time (src/service/service.cpp:1521-1577): a `Request` from the SUB
// socket becomes two tasks (forward, reverse complement), each building and sending one `Reply`.  A GPU index wants
// thousands of queries per launch, so this shell collects the requests of one receive window (a size OR a time limit,
// BatchWindow), answers them with batched calls and emits exactly the replies the reference would have sent:
//
//   CountBatcher   CountReads and ExactMatch + output=count          count_reads, service.cpp:279-315
//   QueryBatcher   ExactMatch + output=reads / all / samples         QueryTask::run + find_reads, :714-797, 1259-1362
//   Worker         the receive loop and its routing                  main(), :1521-1577
//
// Transport stays outside (libzmq is reused, not rebuilt: SURVEY.md section 2 row 20): Worker is a template over a
// type with recv / send / send_count, which INTEGRATION.md binds to the three ZMQ sockets of the reference
// (SUB `pull`, PUSH `push`, PUSH `push_count`); the tests bind it to an in-memory queue.
#ifndef RS_SERVICE_HPP
#define RS_SERVICE_HPP

#include <chrono>
#include <map>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "rs_gpu_bwt.hpp"
#include "rs_wire.hpp"

namespace rsb200 {

// When is a batch due?  When it holds max_pending requests, or when its oldest request has waited `window`.
class BatchWindow {
 public:
  typedef std::chrono::steady_clock clock;
  BatchWindow(size_t max_pending, std::chrono::microseconds window) : max_(max_pending), window_(window), armed_(false) {}
  void on_push(size_t pending_after, clock::time_point now = clock::now()) {
    if (pending_after == 1 || !armed_) {
      first_ = now;
      armed_ = true;
    }
  }
  void on_flush() { armed_ = false; }
  bool due(size_t pending, clock::time_point now = clock::now()) const {
    if (!pending) return false;
    return pending >= max_ || (armed_ && now - first_ >= window_);
  }
  // how long a receiver may block before the batch in hand is due (zero: flush now; window when nothing is pending)
  std::chrono::microseconds wait_budget(size_t pending, clock::time_point now = clock::now()) const {
    if (!pending || !armed_) return window_;
    const std::chrono::microseconds waited = std::chrono::duration_cast<std::chrono::microseconds>(now - first_);
    return waited >= window_ ? std::chrono::microseconds(0) : window_ - waited;
  }

 private:
  size_t max_;
  std::chrono::microseconds window_;
  clock::time_point first_;
  bool armed_;
};

class CountBatcher {
 public:
  explicit CountBatcher(const GpuBWT* bwt, size_t max_pending = (size_t)1 << 20,
                        std::chrono::microseconds window = std::chrono::microseconds(2000))
      : bwt_(bwt), max_pending_(max_pending), window_(max_pending, window) {}

  // One message from the SUB socket.  Returns true when it was a count request (CountReads, or ExactMatch
  // with output=count: the two cases the reference answers with count_reads) and has been queued; false
  // when the caller must route it elsewhere (other request types, or bytes that do not parse).
  bool push(const void* data, size_t size) {
    wire::Request r;
    if (!wire::decode_request(data, size, r)) return false;
    return push(r);
  }
  bool push(const std::string& bytes) { return push(bytes.data(), bytes.size()); }
  bool push(wire::Request& r) {
    int32_t rt;
    if (r.t == wire::CountReads) rt = wire::CountReads;
    else if (r.t == wire::ExactMatch && r.rt == wire::Count) rt = wire::ExactMatch;
    else return false;
    pending_.push_back(Item{std::move(r.q), rt});
    window_.on_push(pending_.size());
    return true;
  }

  size_t pending() const { return pending_.size(); }
  bool full() const { return pending_.size() >= max_pending_; }
  bool due() const { return window_.due(pending_.size()); }
  const BatchWindow& window() const { return window_; }

  // Searches everything queued.  The reference sends the replies of ExactMatch + output=count on its `push` socket
  // (CountTask, service.cpp:1548-1554) and those of CountReads on `push_count` (:1566-1571): send_exact / send_count
  // are called once per Reply accordingly, replies of one sink in arrival order, forward before reverse complement.
  // Returns the number of replies (two per request).
  template <class SendExact, class SendCount>
  size_t flush(SendExact&& send_exact, SendCount&& send_count) {
    const size_t n = pending_.size();
    window_.on_flush();
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
      const std::string a = wire::encode_reply_count(pending_[i].rt, pending_[i].q, false, fwd[i]);
      const std::string b = wire::encode_reply_count(pending_[i].rt, pending_[i].q, true, rev[i]);
      if (pending_[i].rt == wire::CountReads) {
        send_count(a);
        send_count(b);
      } else {
        send_exact(a);
        send_exact(b);
      }
    }
    pending_.clear();
    return 2 * n;
  }
  // one sink for both kinds (a front end that listens on a single channel)
  template <class Send>
  size_t flush(Send&& send) {
    return flush(send, send);
  }

 private:
  struct Item {
    std::string q;
    int32_t rt;
  };
  const GpuBWT* bwt_;
  size_t max_pending_;
  BatchWindow window_;
  std::vector<Item> pending_;
};

// The ExactMatch branch with output = reads / all / samples (QueryTask::run, service.cpp:1259-1362, over find_reads,
// :714-797), per request and strand:
//   query shorter than min_read_length   findInterval of the query as given (find_reads does not screen bytes), the rows
//                                        in the order find_reads emits them (:724-749: the rows past the 2,048-row
//                                        chunks first, then the chunks), each row's read = extractPrefix + extractPostfix
//   min <= length < max_read_length      the query's tiles of min_read_length bases that are reads of this shard
//                                        (suffix filter is_suffix_of, :228-230, then query_exactmatch), then every read
//                                        holding the query (query(), rows in order)
//   length >= max_read_length            tiles of max_read_length, then of min_read_length, that are reads
// For output = all / samples the codes of every read (the RocksDB value, :1323-1347) come from the read-index-addressed
// sample table, names through the hash file's code -> name map; with has_other_meta_data the count and lane bytes
// stored beside each code are sent as c and l (:1338-1342), else 0.  One Reply per strand, forward first, bytes as
// QueryTask serialises them.  Tile matches are emitted in tile order along the query: the reference iterates an
// unordered_set there, so its order among them is unspecified.
// KmerMatch requests (KmerTask, service.cpp:851-990) ride the same batches: the tiles get_tiles(q, k, s) cuts are
// searched as find_reads searches them, their reads united (a set of read strings in the reference: reported once each,
// here in the order the tiles find them), and answered as count (the set's size), reads or all.
class QueryBatcher {
 public:
  QueryBatcher(const GpuBWT* bwt, const std::map<std::string, std::string>* code_to_name, int size_of_sample,
               size_t min_read_length = 73, size_t max_pending = (size_t)1 << 14, uint32_t max_len = 256,
               bool has_other_meta_data = false, const std::string& suffix = std::string(), size_t max_read_length = 100,
               std::chrono::microseconds window = std::chrono::microseconds(2000))
      : bwt_(bwt), hash_(code_to_name), sos_(size_of_sample), min_len_(min_read_length), max_read_(max_read_length),
        max_pending_(max_pending), max_len_(max_len), meta_(has_other_meta_data), suffix_(suffix),
        window_(max_pending, window) {}

  bool push(const void* data, size_t size) {
    wire::Request r;
    if (!wire::decode_request(data, size, r)) return false;
    return push(r);
  }
  bool push(const std::string& bytes) { return push(bytes.data(), bytes.size()); }
  bool push(wire::Request& r) {
    if (r.q.empty()) return false;
    if (r.t == wire::ExactMatch) {
      if (r.rt == wire::Count) return false;  // CountBatcher's
    } else if (r.t == wire::KmerMatch) {
      if (r.k <= 0 || r.s < 0) return false;  // (k = 0 makes get_tiles cut empty tiles: left to the caller)
    } else {
      return false;
    }
    pending_.push_back(Item{std::move(r.q), r.rt, r.t, (size_t)r.k, (size_t)r.s});
    window_.on_push(pending_.size());
    return true;
  }
  size_t pending() const { return pending_.size(); }
  bool full() const { return pending_.size() >= max_pending_; }
  bool due() const { return window_.due(pending_.size()); }
  const BatchWindow& window() const { return window_; }

  static std::string rev_comp(const std::string& w) {  // service.cpp:253-276: other bytes are kept
    std::string s(w.rbegin(), w.rend());
    for (char& c : s) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'C' ? 'G' : c == 'G' ? 'C' : c;
    return s;
  }
  static bool is_suffix_of(const std::string& s, const std::string& w) {  // service.cpp:228-230
    return s.empty() || (w.size() >= s.size() && w.compare(w.size() - s.size(), s.size(), s) == 0);
  }
  static bool acgt(const std::string& w) { return w.find_first_not_of("ACGT") == std::string::npos; }

  template <class Send>
  size_t flush(Send&& send) {
    const size_t n = pending_.size();
    window_.on_flush();
    if (!n) return 0;
    // every string to search, once: the queries themselves (short and middle ones) and the candidate read tiles
    std::vector<std::string> strings;
    struct Piece {  // one contribution to a reply, in order
      size_t string_id;
      int kind;  // 0: rows of a short query in find_reads' order, 1: rows in order (query()), 2: the string itself if a read
    };
    std::vector<std::vector<Piece> > plan(2 * n);
    auto add_tiles = [&](std::vector<Piece>& out, const std::string& w, size_t len) {
      if (w.size() < len) return;
      std::set<std::string> seen;  // get_tiles returns a set (service.cpp:232-246)
      for (size_t i = 0; i + len <= w.size(); ++i) {
        const std::string t = w.substr(i, len);
        if (!seen.insert(t).second) continue;
        if (!is_suffix_of(suffix_, t) || !acgt(t)) continue;  // query_exactmatch is false for other bytes (query.cpp:103)
        strings.push_back(t);
        out.push_back(Piece{strings.size() - 1, 2});
      }
    };
    auto plan_find_reads = [&](std::vector<Piece>& out, const std::string& w) {  // find_reads (service.cpp:714-797)
      const size_t sz = w.size();
      if (sz < min_len_) {
        strings.push_back(w);
        out.push_back(Piece{strings.size() - 1, 0});
      } else if (sz < max_read_) {
        if (sz != min_len_) add_tiles(out, w, min_len_);
        if (acgt(w)) {  // query() answers nothing for other bytes (query.cpp:90)
          strings.push_back(w);
          out.push_back(Piece{strings.size() - 1, 1});
        }
      } else {
        add_tiles(out, w, max_read_);
        if (min_len_ != max_read_) add_tiles(out, w, min_len_);
      }
    };
    for (size_t i = 0; i < n; ++i)
      for (int rev = 0; rev < 2; ++rev) {
        const Item& it = pending_[i];
        const std::string w = rev ? rev_comp(it.q) : it.q;
        std::vector<Piece>& out = plan[2 * i + rev];
        if (it.t == wire::KmerMatch) {
          // find_kmer_reads (service.cpp:466-502): find_reads of every tile get_tiles(w, k, skip) cuts (:232-246) that
          // holds nothing but ACGT (:477); the reads of all tiles are united below
          std::set<std::string> seen;
          for (size_t at = 0; w.size() >= it.k && at + it.k <= w.size(); at += it.s + 1) {
            const std::string tile = w.substr(at, it.k);
            if (!seen.insert(tile).second || !acgt(tile)) continue;
            plan_find_reads(out, tile);
          }
        } else {
          plan_find_reads(out, w);
        }
      }
    const std::vector<BWTInterval> ivs = findIntervalBatch(bwt_, strings);
    // the rows of every string, walked once
    std::vector<uint64_t> rows, first(strings.size() + 1, 0);
    for (size_t j = 0; j < strings.size(); ++j) {
      const BWTInterval& iv = ivs[j];
      if (iv.upper >= iv.lower)
        for (uint64_t r = iv.lower; r <= iv.upper; ++r) rows.push_back(r);
      first[j + 1] = rows.size();
    }
    std::vector<uint64_t> read_index;
    const std::vector<std::string> pre = extractPrefixBatch(bwt_, rows, max_len_);
    const std::vector<std::string> post = extractPostfixBatch(bwt_, rows, max_len_, &read_index);
    // the hits of every reply: position in rows[]
    const uint64_t large = 2048;  // large_match_size, service.cpp:86
    std::vector<std::vector<uint64_t> > hits(2 * n);
    for (size_t j = 0; j < 2 * n; ++j)
      for (const Piece& p : plan[j]) {
        const uint64_t b = first[p.string_id], e = first[p.string_id + 1];
        if (p.kind == 2) {  // query_exactmatch (query.cpp:102-120): some row of the interval rebuilds the string itself
          for (uint64_t h = b; h < e; ++h)
            if (pre[h].empty() && post[h] == strings[p.string_id]) {
              hits[j].push_back(h);
              break;
            }
        } else if (p.kind == 1 || e - b <= 2 * large + 1) {
          for (uint64_t h = b; h < e; ++h) hits[j].push_back(h);
        } else {  // find_reads: the tail past the 2,048-row chunks first, then the chunks (service.cpp:724-749)
          uint64_t start = b;
          while ((e - 1) - start > 2 * large) start += large;
          for (uint64_t h = start; h < e; ++h) hits[j].push_back(h);
          for (uint64_t h = b; h < start; ++h) hits[j].push_back(h);
        }
      }
    // KmerTask keeps a SET of read strings (service.cpp:889): a read found through several tiles is reported once
    for (size_t i = 0; i < n; ++i)
      if (pending_[i].t == wire::KmerMatch)
        for (int rev = 0; rev < 2; ++rev) {
          std::set<std::string> seen;
          std::vector<uint64_t> kept;
          for (uint64_t h : hits[2 * i + rev])
            if (seen.insert(pre[h] + post[h]).second) kept.push_back(h);
          hits[2 * i + rev].swap(kept);
        }
    bool want_samples = false;
    for (size_t i = 0; i < n; ++i) want_samples |= pending_[i].rt != wire::Reads && pending_[i].rt != wire::Count;
    std::vector<uint64_t> soff(rows.size() + 1, 0);
    std::vector<uint8_t> codes, meta;
    if (want_samples && !rows.empty()) {
      codes.resize(1);
      int rc = rs_samples_meta_batch(bwt_->handle(), read_index.data(), rows.size(), soff.data(), 0, codes.data(), nullptr);
      if (rc == RS_STATUS_CAPACITY) {
        codes.resize((size_t)soff[rows.size()] * sos_ + 1);
        if (meta_) meta.resize((size_t)soff[rows.size()] * 2 + 1);
        rc = rs_samples_meta_batch(bwt_->handle(), read_index.data(), rows.size(), soff.data(), soff[rows.size()],
                                   codes.data(), meta_ ? meta.data() : nullptr);
      }
      check(rc);
    }
    for (size_t i = 0; i < n; ++i)
      for (int rev = 0; rev < 2; ++rev) {
        const size_t j = 2 * i + rev;
        const int32_t type = pending_[i].t;
        if (pending_[i].rt == wire::Count) {  // KmerMatch only: the size of the set (service.cpp:893-902)
          send(wire::encode_reply_count(type, pending_[i].q, rev != 0, hits[j].size()));
        } else if (pending_[i].rt == wire::Reads) {
          std::vector<std::string> reads;
          for (uint64_t h : hits[j]) reads.push_back(pre[h] + post[h]);
          send(wire::encode_reply_reads(type, pending_[i].q, rev != 0, reads));
        } else {
          std::vector<wire::ResultAll> out;
          for (uint64_t h : hits[j]) {
            wire::ResultAll one;
            one.r = pre[h] + post[h];
            for (uint64_t c = soff[h]; c < soff[h + 1]; ++c) {
              wire::ReadInfo info;
              const std::string code((const char*)&codes[(size_t)c * sos_], (size_t)sos_);
              const std::map<std::string, std::string>::const_iterator it = hash_->find(code);
              if (it != hash_->end()) info.g = it->second;  // (*hash)[code]: an unknown code reads as ""
              if (meta_) {  // (int)(value[pos]) - 33 on a plain char (service.cpp:1339-1341)
                info.c = (int32_t)(signed char)meta[(size_t)c * 2] - 33;
                info.l = (int32_t)(signed char)meta[(size_t)c * 2 + 1] - 33;
              }
              one.s.push_back(info);
            }
            out.push_back(one);
          }
          send(wire::encode_reply_all(type, pending_[i].q, rev != 0, out, pending_[i].rt));
        }
      }
    pending_.clear();
    return 2 * n;
  }

 private:
  struct Item {
    std::string q;
    int32_t rt, t;  // return type; ExactMatch or KmerMatch
    size_t k, s;    // KmerMatch: tile length and skip
  };
  const GpuBWT* bwt_;
  const std::map<std::string, std::string>* hash_;
  int sos_;
  size_t min_len_, max_read_, max_pending_;
  uint32_t max_len_;
  bool meta_;
  std::string suffix_;
  BatchWindow window_;
  std::vector<Item> pending_;
};

// SiteMatch requests (GtTask::run, service.cpp:993-1212), batched.  Per request and strand: the query, position and
// alternative allele are oriented and the window cut as :1024-1037 does; the windows of a batch go through
// rs_gt_reads_batch (the tile loop over find_gt_reads, :1044-1072) -> read indices -> read strings from the packed read
// table (GpuBWT::loadReadsFile); every read is offered to `align` -- the reference's align_gt_read (ksw.c, host code
// that stays the reference's: bool(read, window, pos, alt, isalt)) -- and the accepted ones are answered as count,
// reads or all (:1121-1199).  The reference keeps the reads in an unordered_set; here they come in read-index order.
// A request whose window the reference itself cannot process (pos beyond the window: it prints an error, :1041-1043;
// a substr past the end or a window shorter than a tile: it throws or never returns) gets the reply head alone.
// Needs the .sai (setSai) and the read table.
template <class Aligner>
class SiteBatcher {
 public:
  SiteBatcher(const GpuBWT* bwt, const std::map<std::string, std::string>* code_to_name, int size_of_sample, Aligner align,
              size_t max_read_length = 100, size_t max_pending = (size_t)1 << 12, uint32_t max_len = 256,
              bool has_other_meta_data = false, uint64_t max_interval_size = 10000,
              std::chrono::microseconds window = std::chrono::microseconds(2000))
      : bwt_(bwt), hash_(code_to_name), sos_(size_of_sample), align_(align), max_read_(max_read_length),
        max_pending_(max_pending), max_len_(max_len), meta_(has_other_meta_data), max_iv_(max_interval_size),
        window_(max_pending, window) {}

  bool push(const std::string& bytes) {
    wire::Request r;
    return wire::decode_request(bytes.data(), bytes.size(), r) && push(r);
  }
  bool push(wire::Request& r) {
    if (r.t != wire::SiteMatch || r.q.empty() || r.k <= 0 || r.s < 0) return false;
    pending_.push_back(Item{std::move(r.q), std::move(r.a), r.rt, (size_t)r.p, (size_t)r.k, (size_t)r.s, r.isalt == 1});
    window_.on_push(pending_.size());
    return true;
  }
  size_t pending() const { return pending_.size(); }
  bool full() const { return pending_.size() >= max_pending_; }
  bool due() const { return window_.due(pending_.size()); }
  const BatchWindow& window() const { return window_; }

  template <class Send>
  size_t flush(Send&& send) {
    const size_t n = pending_.size();
    window_.on_flush();
    if (!n) return 0;
    struct Cut {
      std::string w, alt;
      size_t pos;
      bool bad;
    };
    std::vector<Cut> cuts(2 * n);
    std::map<std::pair<size_t, size_t>, std::vector<size_t> > groups;  // (k, s) -> cuts searched together
    for (size_t i = 0; i < n; ++i)
      for (int rev = 0; rev < 2; ++rev) {
        const Item& it = pending_[i];
        Cut& c = cuts[2 * i + rev];
        c.w = it.q;
        c.alt = it.a;
        c.pos = it.p;
        c.bad = false;
        if (rev) {  // service.cpp:1024-1028
          c.w = QueryBatcher::rev_comp(it.q);
          c.pos = c.w.size() - it.p + 1;
          c.alt = QueryBatcher::rev_comp(it.a);
        }
        if (c.pos <= max_read_) {  // :1030-1036
          c.w = c.w.substr(0, max_read_ + c.pos - 1);
        } else if (c.pos - max_read_ > c.w.size()) {
          c.bad = true;  // substr would throw
        } else {
          c.w = c.w.substr(c.pos - max_read_, max_read_ * 2 - 1);
          c.pos = max_read_;
        }
        if (c.pos > c.w.size() || c.w.size() < it.k) c.bad = true;
        if (!c.bad) groups[std::make_pair(it.k, it.s)].push_back(2 * i + rev);
      }
    std::vector<std::vector<uint64_t> > found(2 * n);
    for (const auto& g : groups) {
      std::vector<std::string> ws;
      std::vector<uint64_t> pos;
      for (size_t j : g.second) {
        ws.push_back(cuts[j].w);
        pos.push_back(cuts[j].pos);
      }
      std::vector<uint32_t> flags;
      const std::vector<std::vector<uint64_t> > ri = findGtReadsBatch(bwt_, ws, pos, g.first.first, g.first.second, max_iv_, &flags);
      for (size_t x = 0; x < g.second.size(); ++x) {
        if (flags[x]) cuts[g.second[x]].bad = true;  // a tile on which the reference throws or hangs
        else found[g.second[x]] = ri[x];
      }
    }
    for (size_t i = 0; i < n; ++i)
      for (int rev = 0; rev < 2; ++rev) {
        const size_t j = 2 * i + rev;
        const Item& it = pending_[i];
        const Cut& c = cuts[j];
        if (c.bad) {
          std::string head;
          wire::reply_head(head, wire::SiteMatch, it.rt, it.q);
          send(head);
          continue;
        }
        const std::vector<std::string> reads = readsBatch(bwt_, found[j], max_len_);
        std::vector<std::string> kept;
        std::vector<uint64_t> kept_index;
        for (size_t h = 0; h < reads.size(); ++h)
          if (align_(reads[h], c.w, c.pos, c.alt, it.isalt)) {
            kept.push_back(reads[h]);
            kept_index.push_back(found[j][h]);
          }
        if (it.rt == wire::Count) {
          send(wire::encode_reply_count(wire::SiteMatch, it.q, rev != 0, kept.size()));
        } else if (it.rt == wire::Reads) {
          send(wire::encode_reply_reads(wire::SiteMatch, it.q, rev != 0, kept));
        } else {
          std::vector<uint64_t> soff(kept.size() + 1, 0);
          std::vector<uint8_t> codes(1), meta;
          if (!kept.empty()) {
            int rc = rs_samples_meta_batch(bwt_->handle(), kept_index.data(), kept.size(), soff.data(), 0, codes.data(), nullptr);
            if (rc == RS_STATUS_CAPACITY) {
              codes.resize((size_t)soff[kept.size()] * sos_ + 1);
              if (meta_) meta.resize((size_t)soff[kept.size()] * 2 + 1);
              rc = rs_samples_meta_batch(bwt_->handle(), kept_index.data(), kept.size(), soff.data(), soff[kept.size()],
                                         codes.data(), meta_ ? meta.data() : nullptr);
            }
            check(rc);
          }
          std::vector<wire::ResultAll> out(kept.size());
          for (size_t h = 0; h < kept.size(); ++h) {
            out[h].r = kept[h];
            for (uint64_t x = soff[h]; x < soff[h + 1]; ++x) {
              wire::ReadInfo info;
              const std::map<std::string, std::string>::const_iterator f =
                  hash_->find(std::string((const char*)&codes[(size_t)x * sos_], (size_t)sos_));
              if (f != hash_->end()) info.g = f->second;
              if (meta_) {
                info.c = (int32_t)(signed char)meta[(size_t)x * 2] - 33;
                info.l = (int32_t)(signed char)meta[(size_t)x * 2 + 1] - 33;
              }
              out[h].s.push_back(info);
            }
          }
          send(wire::encode_reply_all(wire::SiteMatch, it.q, rev != 0, out, it.rt));
        }
      }
    pending_.clear();
    return 2 * n;
  }

 private:
  struct Item {
    std::string q, a;
    int32_t rt;
    size_t p, k, s;
    bool isalt;
  };
  const GpuBWT* bwt_;
  const std::map<std::string, std::string>* hash_;
  int sos_;
  Aligner align_;
  size_t max_read_, max_pending_;
  uint32_t max_len_;
  bool meta_;
  uint64_t max_iv_;
  BatchWindow window_;
  std::vector<Item> pending_;
};

// The worker's receive loop (main(), service.cpp:1521-1577) over any transport:
//
//   struct Transport {
//     bool recv(std::string& bytes, std::chrono::microseconds wait);   // false: nothing arrived within `wait`
//     void send(const std::string& reply);                             // the PUSH socket `push`
//     void send_count(const std::string& reply);                       // the PUSH socket `push_count`
//     bool stopped();                                                  // the reference loops forever
//   };
//
// Count requests, ExactMatch and KmerMatch queries are batched (windows above) and answered on the GPU; SiteMatch
// requests -- whose seed stage is rs_gt_reads_batch and whose reply assembly needs the Smith-Waterman aligner that
// stays on the host -- are handed to `other`, as are bytes that do not parse (the reference prints "unrecognised
// request type").
template <class Transport, class Other>
class Worker {
 public:
  Worker(Transport* io, CountBatcher* counts, QueryBatcher* queries, Other other)
      : io_(io), counts_(counts), queries_(queries), other_(other), served_(0) {}

  // one pass: drain what has arrived (without waiting longer than the batches in hand allow), flush what is due
  void poll() {
    for (;;) {
      std::chrono::microseconds wait = counts_->window().wait_budget(counts_->pending());
      const std::chrono::microseconds w2 = queries_->window().wait_budget(queries_->pending());
      if (w2 < wait) wait = w2;
      std::string bytes;
      if (wait.count() == 0 || !io_->recv(bytes, wait)) break;
      wire::Request r;
      if (!wire::decode_request(bytes.data(), bytes.size(), r)) {
        other_(bytes);
        continue;
      }
      if (!counts_->push(r) && !queries_->push(r)) other_(bytes);
      if (counts_->full() || queries_->full()) break;
    }
    flush_due(false);
  }
  void run() {
    while (!io_->stopped()) poll();
    flush_due(true);
  }
  void flush_due(bool all) {
    Transport* io = io_;
    if (all || counts_->due())
      served_ += counts_->flush([io](const std::string& b) { io->send(b); }, [io](const std::string& b) { io->send_count(b); });
    if (all || queries_->due()) served_ += queries_->flush([io](const std::string& b) { io->send(b); });
  }
  size_t replies_sent() const { return served_; }

 private:
  Transport* io_;
  CountBatcher* counts_;
  QueryBatcher* queries_;
  Other other_;
  size_t served_;
};

}  // namespace rsb200

#endif
