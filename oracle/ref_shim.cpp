// TEST INFRASTRUCTURE ONLY -- never linked into or called by the product path.
//
// extern "C" shim over the UNMODIFIED reference library (wtsi-svi/ReadServer
// src/bwt/{query,rlebwt,rlebwt_reader}.cpp, include/bwt/*.h).  The reference
// sources are compiled where they lie under /root/reference by oracle/Makefile
// and linked with this file into oracle/_ref/libref_bwt.so.  Nothing here
// re-implements reference arithmetic: every function forwards to the public
// `class BWT` interface (include/bwt/bwt.h:6-15) or to the free functions of
// include/bwt/query.h:18-32.  The only loops are batch/thread drivers and the
// two "terminal row" walkers, which replay extractPrefix / extractPostfix
// (src/bwt/query.cpp:43-85) through the same public getters so that the
// numeric read identity can be observed.
#include <stdint.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>
#include <chrono>

#include "rlebwt.h"
#include "query.h"

namespace {
struct RefHandle {
  RLEBWT* bwt;
};

template <class F>
void run_threads(size_t n, int nthreads, F f) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads == 1 || n < 2) {
    f(0, n, 1);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back([=]() { f((size_t)t, n, (size_t)nthreads); });
  for (auto& t : th) t.join();
}
}  // namespace

extern "C" {

// RLEBWT(filename) -- src/bwt/rlebwt.cpp:11-32 (loads <path>.bpi2 if present, else builds the tree)
void* ref_open(const char* bwt_path) {
  RefHandle* h = new RefHandle;
  h->bwt = new RLEBWT(std::string(bwt_path));
  return h;
}

void ref_close(void* hv) {
  RefHandle* h = (RefHandle*)hv;
  if (!h) return;
  delete h->bwt;
  delete h;
}

// src/util/index_rlebwt.cpp:22-23
void ref_serialise(void* hv, const char* bpi_path) {
  ((RefHandle*)hv)->bwt->serialiseFMIndex(std::string(bpi_path));
}

uint64_t ref_bwlen(void* hv) { return ((RefHandle*)hv)->bwt->getBWLen(); }
uint64_t ref_pc(void* hv, char b) { return ((RefHandle*)hv)->bwt->getPC(b); }
uint64_t ref_get_occ(void* hv, char b, uint64_t i) { return ((RefHandle*)hv)->bwt->getOcc(b, i); }
uint64_t ref_get_occ_at(void* hv, char b, uint64_t c) { return ((RefHandle*)hv)->bwt->getOccAt(b, c); }
char ref_get_char(void* hv, uint64_t i) { return ((RefHandle*)hv)->bwt->getChar(i); }
char ref_get_f(void* hv, uint64_t i) { return ((RefHandle*)hv)->bwt->getF(i); }

void ref_get_occ_batch(void* hv, const char* b, const uint64_t* idx, size_t n, uint64_t* out) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  for (size_t i = 0; i < n; ++i) out[i] = p->getOcc(b[i], idx[i]);
}

void ref_get_char_batch(void* hv, const uint64_t* idx, size_t n, char* out) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  for (size_t i = 0; i < n; ++i) out[i] = p->getChar(idx[i]);
}

void ref_get_occ_at_batch(void* hv, const char* b, const uint64_t* cnt, size_t n, uint64_t* out) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  for (size_t i = 0; i < n; ++i) out[i] = p->getOccAt(b[i], cnt[i]);
}

// findInterval (src/bwt/query.cpp:24-41) over Q fixed-length ASCII k-mers, striding threads.
// Returns wall seconds spent inside the query loop (steady_clock), as BASELINE.md section 3 asks.
double ref_find_interval_batch(void* hv, const char* kmers, int k, size_t Q, uint64_t* lower,
                               uint64_t* upper, int nthreads) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  auto t0 = std::chrono::steady_clock::now();
  run_threads(Q, nthreads, [=](size_t t, size_t n, size_t stride) {
    std::string w;
    for (size_t q = t; q < n; q += stride) {
      w.assign(kmers + q * (size_t)k, (size_t)k);
      BWTInterval iv = findInterval(p, w);
      lower[q] = iv.lower;
      upper[q] = iv.upper;
    }
  });
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// extractPrefix + extractPostfix (src/bwt/query.cpp:43-85) for one row; returns the string length
// (truncated to cap-1 chars in `out`, always NUL terminated) and the prefix length in *prefix_len.
size_t ref_extract_read(void* hv, uint64_t row, char* out, size_t cap, size_t* prefix_len) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  size_t r = (size_t)row;
  std::string pre = extractPrefix(p, r);
  std::string post = extractPostfix(p, r);
  if (prefix_len) *prefix_len = pre.size();
  std::string s = pre + post;
  size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
  memcpy(out, s.data(), n);
  out[n] = 0;
  return s.size();
}

// Replay of the extractPrefix loop (src/bwt/query.cpp:43-63) through the public getters, keeping the
// terminal idx (= lexicographic rank of the read) and the number of bases walked.
// A walk that enters the reference's tail-clamp region can cycle forever -- the
// reference's extractPrefix would hang; the replay gives up after max_steps bases and reports
// idx = UINT64_MAX, steps = UINT32_MAX.
static inline uint64_t prefix_terminal(const BWT* p, uint64_t row, uint32_t* steps, uint32_t max_steps) {
  uint64_t idx = row;
  uint32_t s = 0;
  while (1) {
    char b = p->getChar(idx);
    idx = p->getPC(b) + p->getOcc(b, idx - 1);
    if (b == '$') break;
    if (++s > max_steps) { s = UINT32_MAX; idx = UINT64_MAX; break; }
  }
  if (steps) *steps = s;
  return idx;
}

// Replay of the extractPostfix loop (src/bwt/query.cpp:65-85); reports the row (< num_strings) at which
// getF returned '$' = the read's input-order index (the select of that last step is still performed).
static inline uint64_t postfix_terminal(const BWT* p, uint64_t row, uint32_t* steps, uint32_t max_steps) {
  uint64_t idx = row;
  uint32_t s = 0;
  while (1) {
    char f = p->getF(idx);
    uint64_t fc = idx - p->getPC(f) + 1;
    const uint64_t at = idx;  // row in F: when f == '$' it is the read's input-order index
    idx = p->getOccAt(f, fc);
    if (f == '$') { idx = at; break; }
    if (++s > max_steps) { s = UINT32_MAX; idx = UINT64_MAX; break; }
  }
  if (steps) *steps = s;
  return idx;
}

double ref_locate_batch(void* hv, const uint64_t* rows, size_t n, uint64_t* lexrank, uint32_t* pre_steps,
                        uint64_t* read_index, uint32_t* post_steps, int nthreads, uint32_t max_steps) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  auto t0 = std::chrono::steady_clock::now();
  run_threads(n, nthreads, [=](size_t t, size_t nn, size_t stride) {
    for (size_t i = t; i < nn; i += stride) {
      uint32_t s1 = 0, s2 = 0;
      if (lexrank) lexrank[i] = prefix_terminal(p, rows[i], &s1, max_steps);
      if (read_index) read_index[i] = postfix_terminal(p, rows[i], &s2, max_steps);
      if (pre_steps) pre_steps[i] = s1;
      if (post_steps) post_steps[i] = s2;
    }
  });
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// Timed extractPrefix+extractPostfix (the reference's own string path) over rows; returns seconds and
// a checksum of the total bases produced so the work cannot be optimised away.
double ref_extract_batch_timed(void* hv, const uint64_t* rows, size_t n, int nthreads, uint64_t* total_bases) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  std::vector<uint64_t> acc((size_t)(nthreads < 1 ? 1 : nthreads), 0);
  uint64_t* accp = acc.data();
  auto t0 = std::chrono::steady_clock::now();
  run_threads(n, nthreads, [=](size_t t, size_t nn, size_t stride) {
    uint64_t a = 0;
    for (size_t i = t; i < nn; i += stride) {
      size_t r = (size_t)rows[i];
      a += extractPrefix(p, r).size() + extractPostfix(p, r).size();
    }
    accp[t] = a;
  });
  auto t1 = std::chrono::steady_clock::now();
  uint64_t s = 0;
  for (auto v : acc) s += v;
  if (total_bases) *total_bases = s;
  return std::chrono::duration<double>(t1 - t0).count();
}

// query / query_exactmatch (src/bwt/query.cpp:87-120) -- returns number of reads; strings joined by '\n'.
size_t ref_query(void* hv, const char* w, char* out, size_t cap) {
  const BWT* p = ((RefHandle*)hv)->bwt;
  std::vector<std::string> v = query(p, std::string(w));
  size_t pos = 0;
  for (auto& s : v) {
    if (pos + s.size() + 1 >= cap) break;
    memcpy(out + pos, s.data(), s.size());
    pos += s.size();
    out[pos++] = '\n';
  }
  if (cap) out[pos < cap ? pos : cap - 1] = 0;
  return v.size();
}

int ref_query_exactmatch(void* hv, const char* w) {
  return query_exactmatch(((RefHandle*)hv)->bwt, std::string(w)) ? 1 : 0;
}

}  // extern "C"
