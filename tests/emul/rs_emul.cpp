// TEST INFRASTRUCTURE ONLY -- never part of the product library.
//
// Host replay of the device code: the REAL per-query operations (readserver_b200/csrc/rs_ops.h, the
// source the sm_100a kernels are compiled from) run here on host memory over the REAL host builder
// (rs_build.cpp) -- the kernel-side view's pointers simply point into the builder's arrays.  It lets
// the CPU test suite (-m "not gpu") check the HBM layout, the directory logic, the in-block
// rank/access/select arithmetic and the reference-tail emulation against the oracle without a GPU.
// The product never loads this file: librs_b200.so has no CPU compute path.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "rs_build.h"
#include "rs_layout.h"
#include "rs_gt_ops.h"
#include "rs_ops.h"

namespace {
struct Emul {
  RsHostIndex h;
  rs_index_dev v;
};

inline uint32_t code_of(uint8_t b) { return b == 'A' ? 1u : b == 'C' ? 2u : b == 'G' ? 3u : b == 'T' ? 4u : 0u; }
}  // namespace

extern "C" {

void* emul_open(const uint8_t* runs, uint64_t nruns, uint64_t num_strings, uint64_t num_symbols, int* status) {
  Emul* e = new Emul;
  const int rc = rs_build_host_index(runs, nruns, num_strings, num_symbols, e->h);
  if (status) *status = rc;
  if (rc != RS_OK) {
    delete e;
    return nullptr;
  }
  rs_fill_view(e->h, &e->v);
  e->v.blocks = e->h.blocks.data();
  e->v.dir = e->h.dir.data();
  for (int c = 0; c < RS_NSYM; ++c) e->v.sel[c] = e->h.sel[c].data();
  return e;
}

void* emul_open_file(const char* path, int* status) {
  std::vector<uint8_t> runs;
  uint64_t ns = 0, nsym = 0;
  int rc = rs_read_bwt_file(path, runs, ns, nsym);
  if (rc != RS_OK) {
    if (status) *status = rc;
    return nullptr;
  }
  return emul_open(runs.data(), runs.size(), ns, nsym, status);
}

void emul_close(void* p) { delete (Emul*)p; }
void emul_set_tail(void* p, int on) {
  Emul* e = (Emul*)p;
  rs_view_set_tail(e->h, on != 0, &e->v);
}

/* out[0..] = n, nblocks, ndir, clamp_c, vlast, depth, q2_safe, pc[5], tot[5], init_occ[5], clamp_char */
void emul_info(void* p, uint64_t* out) {
  const RsHostIndex& h = ((Emul*)p)->h;
  out[0] = h.n; out[1] = h.blocks.size(); out[2] = h.dir.size(); out[3] = h.clamp_c; out[4] = h.vlast;
  out[5] = h.depth; out[6] = h.q2_safe; out[22] = h.clamp_char;
  for (int c = 0; c < 5; ++c) { out[7 + c] = h.pc[c]; out[12 + c] = h.tot[c]; out[17 + c] = h.init_occ[c]; }
}

/* rs_parse_etl_samples: returns status; *nreads, *ncodes sizes; data copied out by a second call */
int emul_parse_etl(const char* path, int sos, int meta, uint64_t* offsets, uint64_t cap_off, uint8_t* codes, uint64_t cap_codes,
                   uint64_t* nreads, uint64_t* ncodes) {
  std::vector<uint64_t> off;
  std::vector<uint8_t> cd;
  const int rc = rs_parse_etl_samples(path, sos, meta, off, cd);
  *nreads = off.size() - 1;
  *ncodes = cd.size();
  if (rc == RS_OK && offsets && off.size() <= cap_off && cd.size() <= cap_codes) {
    memcpy(offsets, off.data(), off.size() * 8);
    if (!cd.empty()) memcpy(codes, cd.data(), cd.size());
  }
  return rc;
}

/* raw 64-byte blocks, for the layout tests */
const void* emul_blocks(void* p) { return ((Emul*)p)->h.blocks.data(); }

void emul_occ_batch(void* p, const uint8_t* b, const uint64_t* idx, uint64_t n, uint64_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) out[i] = rs_occ_byte(e.v, b[i], idx[i]);
}

void emul_char_batch(void* p, const uint64_t* idx, uint64_t n, uint8_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t s = rs_char(e.v, idx[i]);
    out[i] = s <= 4 ? (uint8_t)"$ACGT"[s] : (uint8_t)'?';
  }
}

void emul_occ_at_batch(void* p, const uint8_t* b, const uint64_t* cnt, uint64_t n, uint64_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) out[i] = rs_occ_at(e.v, code_of(b[i]), cnt[i]);
}

void emul_find_interval_batch(void* p, const uint8_t* kmers, int k, uint64_t Q, uint64_t* lower, uint64_t* upper) {
  const Emul& e = *(Emul*)p;
  /* the per-string search K1b and the /gt seed kernel run (any byte value); K1's own loop spreads the same
   * rs_occ2 scans over one or two trips */
  for (uint64_t q = 0; q < Q; ++q) rs_find_interval_one(e.v, kmers + q * (uint64_t)k, (uint64_t)k, lower + q, upper + q);
}

void emul_prefix_walk_batch(void* p, const uint64_t* rows, uint64_t n, uint32_t max_steps, uint64_t* lexrank,
                            uint32_t* steps) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t idx = rows[i];
    uint32_t s = 0;
    while (true) {
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(e.v, idx, &nxt);
      idx = nxt;
      if (sym == 0) break;
      if (++s > max_steps) {
        s = 0xFFFFFFFFu;
        idx = ~0ull;
        break;
      }
    }
    lexrank[i] = idx;
    steps[i] = s;
  }
}

void emul_postfix_walk_batch(void* p, const uint64_t* rows, uint64_t n, uint32_t max_steps, uint64_t* read_index,
                             uint32_t* steps) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t idx = rows[i], term = ~0ull;
    uint32_t s = 0;
    while (true) {
      uint64_t nxt;
      const uint64_t at = idx;
      const uint32_t f = rs_fl_step(e.v, idx < e.v.n ? idx : e.v.n - 1, &nxt);
      idx = nxt;
      if (f == 0) {
        term = at;
        break;
      }
      if (++s > max_steps || idx == ~0ull) {
        s = 0xFFFFFFFFu;
        break;
      }
    }
    read_index[i] = term;
    steps[i] = s;
  }
}

/* the per-tile seed logic of the /gt seed stage (rs_gt_ops.h, the source k_gt_seeds runs): out[i] =
 * {lo0, hi0, thr0, lo1, hi1, thr1, kind, flags} for tile i of window w */
void emul_gt_seeds(void* p, const uint8_t* w, uint64_t L, uint64_t pos, uint64_t ntiles, uint32_t kmers, uint32_t step,
                   uint64_t max_iv, uint64_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < ntiles; ++i) {
    gt_iv seed[2];
    uint64_t thr[2];
    uint32_t kind, flag;
    rs_gt_tile_seeds(e.v, w, L, pos, i, kmers, step, max_iv, seed, thr, &kind, &flag);
    uint64_t* o = out + 8 * i;
    o[0] = seed[0].lo, o[1] = seed[0].hi, o[2] = thr[0];
    o[3] = seed[1].lo, o[4] = seed[1].hi, o[5] = thr[1];
    o[6] = kind, o[7] = flag;
  }
}

}  // extern "C"
