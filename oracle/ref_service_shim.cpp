// TEST INFRASTRUCTURE ONLY.
// extern "C" wrapper around functions of the reference's query worker that are cut out of
// /root/reference/src/service/service.cpp AT BUILD TIME by oracle/Makefile (awk, by function signature) into
// oracle/_ref/service_extract.inc -- is_suffix_of, get_tiles, rev_comp, find_gt_reads (service.cpp:227-276, 507-711),
// unmodified.  service.cpp as a whole needs ZeroMQ, protobuf, RocksDB and libconfig, none of which are in this image;
// these four functions need only <string>, <unordered_set> and the reference's query.h.  Nothing of the reference's
// source is copied into this repository: the extract lives in oracle/_ref/ (git-ignored).
//
// The one liberty taken: the reference's constant `max_interval_size` (service.cpp:91, 10000) is made a variable so
// that tests can drive find_gt_reads' interval-extension loops on small BWTs.
#include <string.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <unordered_set>
#include <vector>

#include "bwt.h"
#include "query.h"
#include "rlebwt.h"

using namespace std;

static size_t g_max_interval_size = 10000;  // service.cpp:91
#define max_interval_size g_max_interval_size
static size_t MAX_READ_LENGTH = 100;        // service.cpp:56
static size_t MIN_READ_LENGTH = 73;         // service.cpp:57

#include "service_extract.inc"

static size_t join(const unordered_set<string>& s, char* out, size_t cap) {
  vector<string> v(s.begin(), s.end());
  sort(v.begin(), v.end());
  size_t pos = 0;
  for (auto& x : v) {
    if (pos + x.size() + 1 >= cap) return (size_t)-2;  // buffer too small
    memcpy(out + pos, x.data(), x.size());
    pos += x.size();
    out[pos++] = '\n';
  }
  if (cap) out[pos < cap ? pos : cap - 1] = 0;
  return v.size();
}

extern "C" {

void* refsvc_open(const char* bwt_path) { return new RLEBWT(string(bwt_path)); }
void refsvc_close(void* h) { delete (RLEBWT*)h; }
void refsvc_set_limits(size_t max_interval, size_t max_read_length, size_t min_read_length) {
  g_max_interval_size = max_interval;
  MAX_READ_LENGTH = max_read_length;
  MIN_READ_LENGTH = min_read_length;
}

// get_tiles (service.cpp:232-246): tiles sorted, '\n' separated; returns their number (the set drops duplicates)
size_t refsvc_get_tiles(const char* w, size_t kmer, size_t skip, char* out, size_t cap) {
  return join(get_tiles(string(w), kmer, skip), out, cap);
}

int refsvc_is_suffix_of(const char* s, const char* w) { return is_suffix_of(string(s), string(w)) ? 1 : 0; }

size_t refsvc_rev_comp(const char* w, char* out, size_t cap) {
  const string r = rev_comp(string(w));
  if (r.size() + 1 > cap) return (size_t)-2;
  memcpy(out, r.c_str(), r.size() + 1);
  return r.size();
}

// find_gt_reads (service.cpp:507-711) for tile `index` of window w: the reads, sorted, '\n' separated.
// Returns their number, (size_t)-1 when the reference throws (std::out_of_range from substr), -2: buffer too small.
size_t refsvc_find_gt_reads(void* h, const char* w, size_t index, size_t kmers, size_t skip, size_t pos, char* out,
                            size_t cap) {
  try {
    return join(find_gt_reads((const BWT*)(RLEBWT*)h, string(w), index, kmers, skip, pos), out, cap);
  } catch (const std::out_of_range&) {
    return (size_t)-1;
  }
}

// The tile loop of GtTask::run (service.cpp:1044-1072) over an already cut window: union over tiles 0 .. tiles-1.
size_t refsvc_gt_task_reads(void* h, const char* w, size_t kmers, size_t skip, size_t pos, char* out, size_t cap) {
  const string q(w);
  if (pos > q.size() || q.size() < kmers) return 0;
  const size_t tiles = ((q.size() - kmers) / (skip + 1)) + 1;
  unordered_set<string> seqs;
  try {
    for (size_t i = 0; i < tiles; ++i) {
      const unordered_set<string> v = find_gt_reads((const BWT*)(RLEBWT*)h, q, i, kmers, skip, pos);
      seqs.insert(v.begin(), v.end());
    }
  } catch (const std::out_of_range&) {
    return (size_t)-1;
  }
  return join(seqs, out, cap);
}

}  // extern "C"
