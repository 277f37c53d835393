// TEST INFRASTRUCTURE ONLY -- never part of the product library.
//
// Lane-by-lane host replay of the device lane-group code (readserver_b200/csrc/rs_device.cuh) on top
// of the REAL host builder (rs_build.cpp) and the REAL per-lane SWAR arithmetic (rs_swar.h): the
// 1, 2, 4 or 8 lanes of a group are looped over and warp shuffles become array reads.  It lets the CPU test
// suite (-m "not gpu") check the HBM layout, the directory logic, the in-block rank/access/select
// arithmetic and the reference-tail emulation against the oracle without a GPU.  The product never
// loads this file: librs_b200.so has no CPU compute path.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "rs_build.h"
#include "rs_layout.h"
#include "rs_swar.h"

namespace {
struct Emul {
  RsHostIndex h;
  bool emulate_tail = true;
  int lanes = 8;
};

constexpr int HDR_WORDS = RS_HDR_BYTES / 4;

/* one block split over LANES lanes, as rs_blk_words does on the device */
template <int LANES>
struct Hit {
  uint32_t w[LANES][32 / LANES];
  uint32_t lane_base[LANES];
  uint32_t off;
  uint64_t start;
  const rs_block* blk;
};

inline uint64_t ref_pos(const Emul& e, uint64_t i) {
  if (e.emulate_tail && ((i + 1) >> 16) >= e.h.vlast && i >= e.h.clamp_c) i = e.h.clamp_c - 1;
  return i;
}

inline uint64_t ref_pos_char(const Emul& e, uint64_t i) {
  if (e.emulate_tail && ((i + 1) >> 16) >= e.h.vlast && i >= e.h.clamp_char) i = e.h.clamp_char - 1;
  return i;
}

uint64_t dir_block(uint64_t e, uint32_t o, bool* more) {
  const uint32_t b1 = (uint32_t)(e >> 32) & RS_SLOT_MASK;
  const uint32_t b2 = (uint32_t)(e >> 42) & RS_SLOT_MASK;
  const uint32_t b3 = (uint32_t)(e >> 52) & RS_SLOT_MASK;
  const uint32_t add = (uint32_t)(b1 != 0 && o >= b1) + (uint32_t)(b2 != 0 && o >= b2) + (uint32_t)(b3 != 0 && o >= b3);
  *more = (e & RS_DIR_MORE) != 0 && add == 3;
  return (uint64_t)(uint32_t)e + add;
}

template <int LANES>
uint32_t split_block(const rs_block* blk, Hit<LANES>* H) {
  constexpr int NW = 32 / LANES;
  uint32_t words[32];
  memcpy(words, blk, RS_BLOCK_BYTES);
  uint32_t run = 0;
  for (int g = 0; g < LANES; ++g) {
    for (int j = 0; j < NW; ++j) H->w[g][j] = (g * NW + j >= HDR_WORDS) ? words[g * NW + j] : 0u;
    H->lane_base[g] = run;
    run += rs_words_total<NW>(H->w[g]);
  }
  H->start = blk->start;
  H->blk = blk;
  return run;
}

template <int LANES>
void fetch(const Emul& e, uint64_t pos, Hit<LANES>* H) {
  bool more;
  uint64_t b = dir_block(e.h.dir[pos >> RS_SLOT_SHIFT], (uint32_t)pos & RS_SLOT_MASK, &more);
  while (true) {
    const uint32_t span = split_block(&e.h.blocks[b], H);
    const uint64_t off = pos - H->start + 1;
    if (more && off > span) {
      ++b;
      continue;
    }
    H->off = (uint32_t)off;
    return;
  }
}

inline uint64_t count_before(const rs_block* blk, uint32_t c) {
  if (c) return blk->cnt[c - 1];
  return blk->start - (blk->cnt[0] + blk->cnt[1] + blk->cnt[2] + blk->cnt[3]);
}

template <int LANES>
uint64_t hit_rank(const Hit<LANES>& H, uint32_t c) {
  uint32_t sum = 0;
  for (int g = 0; g < LANES; ++g) {
    uint32_t part, unused;
    rs_words_rank<32 / LANES, false>(H.w[g], c, (int32_t)H.off - (int32_t)H.lane_base[g], 0, &part, &unused);
    sum += part;
  }
  return count_before(H.blk, c) + sum;
}

template <int LANES>
uint32_t hit_symbol(const Hit<LANES>& H) {
  uint32_t sym = 0xFF;
  int owners = 0;
  for (int g = 0; g < LANES; ++g) {
    const uint32_t s = rs_words_symbol<32 / LANES>(H.w[g], (int32_t)H.off - (int32_t)H.lane_base[g]);
    if (s != 0xFF) {
      sym = s;
      ++owners;
    }
  }
  return owners == 1 ? sym : 0xFE; /* exactly one lane must own the position */
}

template <int LANES>
uint64_t occ_t(const Emul& e, uint32_t c, uint64_t i) {
  if (i == ~0ull) return 0;
  uint64_t pos = ref_pos(e, i);
  if (pos >= e.h.n) pos = e.h.n - 1;
  Hit<LANES> H;
  fetch(e, pos, &H);
  return hit_rank(H, c);
}

/* rs_grp_occ2: both ranks from one scan when the positions share a block */
template <int LANES>
void occ2_t(const Emul& e, uint32_t c, uint64_t i1, uint64_t i2, uint64_t* r1, uint64_t* r2) {
  const bool none1 = i1 == ~0ull, none2 = i2 == ~0ull;
  uint64_t p1 = none1 ? 0 : ref_pos(e, i1), p2 = none2 ? 0 : ref_pos(e, i2);
  if (p1 >= e.h.n) p1 = e.h.n - 1;
  if (p2 >= e.h.n) p2 = e.h.n - 1;
  bool m1, m2;
  const uint64_t b1 = dir_block(e.h.dir[p1 >> RS_SLOT_SHIFT], (uint32_t)p1 & RS_SLOT_MASK, &m1);
  const uint64_t b2 = dir_block(e.h.dir[p2 >> RS_SLOT_SHIFT], (uint32_t)p2 & RS_SLOT_MASK, &m2);
  if (b1 != b2 || m1 || m2 || p2 < p1) {
    *r1 = occ_t<LANES>(e, c, i1);
    *r2 = occ_t<LANES>(e, c, i2);
    return;
  }
  Hit<LANES> H;
  split_block(&e.h.blocks[b1], &H);
  const uint32_t off = (uint32_t)(p1 - H.start + 1);
  uint32_t s1 = 0, s2 = 0;
  for (int g = 0; g < LANES; ++g) {
    const int32_t t1 = (int32_t)off - (int32_t)H.lane_base[g];
    const int32_t t2 = t1 + (int32_t)(uint32_t)(p2 - p1);
    uint32_t a1, a2;
    rs_words_rank<32 / LANES, true>(H.w[g], c, t1, t2, &a1, &a2);
    s1 += a1;
    s2 += a2;
  }
  const uint64_t before = count_before(H.blk, c);
  *r1 = none1 ? 0 : before + s1;
  *r2 = none2 ? 0 : before + s2;
}

template <int LANES>
uint32_t chr_t(const Emul& e, uint64_t i) {
  uint64_t pos = ref_pos_char(e, i);
  if (pos >= e.h.n) pos = e.h.n - 1;
  Hit<LANES> H;
  fetch(e, pos, &H);
  return hit_symbol(H);
}

template <int LANES>
uint32_t lf_step_t(const Emul& e, uint64_t idx, uint64_t* next) {
  uint64_t pc = ref_pos_char(e, idx);
  if (pc >= e.h.n) pc = e.h.n - 1;
  Hit<LANES> H;
  fetch(e, pc, &H);
  uint32_t sym = hit_symbol(H);
  if (sym > 4) sym = 0;
  const bool none = idx == 0;
  const uint64_t pp = none ? 0 : ref_pos(e, idx - 1);
  uint64_t o;
  if (!none && pp + 1 == pc)
    o = hit_rank(H, sym) - 1;
  else
    o = occ_t<LANES>(e, sym, idx - 1);
  *next = e.h.pc[sym] + o;
  return sym;
}

/* select: position of the r-th (1-based) occurrence of c */
template <int LANES>
uint64_t occ_at_t(const Emul& e, uint32_t c, uint64_t r) {
  constexpr int NW = 32 / LANES;
  const RsHostIndex& h = e.h;
  if (r == 0 || r > h.tot[c]) return ~0ull;
  uint64_t b = h.sel[c][(r - 1) >> h.sel_shift[c]];
  while (true) {
    Hit<LANES> H;
    split_block(&h.blocks[b], &H);
    uint32_t cbase[LANES], crun = 0;
    for (int g = 0; g < LANES; ++g) {
      cbase[g] = crun;
      crun += rs_words_count_all<NW>(H.w[g], c);
    }
    const uint64_t want = r - count_before(H.blk, c);
    if (want > crun) {
      ++b;
      continue;
    }
    for (int g = 0; g < LANES; ++g) {
      const uint32_t p = rs_words_select<NW>(H.w[g], c, (int32_t)want - (int32_t)cbase[g]);
      if (p) return H.start + H.lane_base[g] + p - 1;
    }
    return ~0ull - 1;
  }
}

#define BY_LANES(e, CALL)                 \
  switch ((e).lanes) {                    \
    case 1: return CALL(1);               \
    case 2: return CALL(2);               \
    case 4: return CALL(4);               \
    default: return CALL(8);              \
  }

uint64_t occ(const Emul& e, uint32_t c, uint64_t i) {
#define CALL(L) occ_t<L>(e, c, i)
  BY_LANES(e, CALL)
#undef CALL
}
void occ2(const Emul& e, uint32_t c, uint64_t i1, uint64_t i2, uint64_t* r1, uint64_t* r2) {
#define CALL(L) occ2_t<L>(e, c, i1, i2, r1, r2)
  BY_LANES(e, CALL)
#undef CALL
}
uint32_t chr(const Emul& e, uint64_t i) {
#define CALL(L) chr_t<L>(e, i)
  BY_LANES(e, CALL)
#undef CALL
}
uint32_t lf_step(const Emul& e, uint64_t idx, uint64_t* next) {
#define CALL(L) lf_step_t<L>(e, idx, next)
  BY_LANES(e, CALL)
#undef CALL
}
uint64_t occ_at(const Emul& e, uint32_t c, uint64_t r) {
#define CALL(L) occ_at_t<L>(e, c, r)
  BY_LANES(e, CALL)
#undef CALL
}

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
void emul_set_tail(void* p, int on) { ((Emul*)p)->emulate_tail = on != 0; }
void emul_set_lanes(void* p, int lanes) { ((Emul*)p)->lanes = lanes; }

/* out[0..] = n, nblocks, ndir, clamp_c, vlast, depth, q2_safe, pc[5], tot[5], init_occ[5], clamp_char */
void emul_info(void* p, uint64_t* out) {
  const RsHostIndex& h = ((Emul*)p)->h;
  out[0] = h.n; out[1] = h.blocks.size(); out[2] = h.dir.size(); out[3] = h.clamp_c; out[4] = h.vlast;
  out[5] = h.depth; out[6] = h.q2_safe; out[22] = h.clamp_char;
  for (int c = 0; c < 5; ++c) { out[7 + c] = h.pc[c]; out[12 + c] = h.tot[c]; out[17 + c] = h.init_occ[c]; }
}

void emul_occ_batch(void* p, const uint8_t* b, const uint64_t* idx, uint64_t n, uint64_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) out[i] = occ(e, code_of(b[i]), idx[i]);
}

void emul_char_batch(void* p, const uint64_t* idx, uint64_t n, uint8_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t s = chr(e, idx[i]);
    out[i] = s <= 4 ? (uint8_t)"$ACGT"[s] : (uint8_t)'?';
  }
}

void emul_occ_at_batch(void* p, const uint8_t* b, const uint64_t* cnt, uint64_t n, uint64_t* out) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) out[i] = occ_at(e, code_of(b[i]), cnt[i]);
}

void emul_find_interval_batch(void* p, const uint8_t* kmers, int k, uint64_t Q, uint64_t* lower, uint64_t* upper) {
  const Emul& e = *(Emul*)p;
  for (uint64_t q = 0; q < Q; ++q) {
    const uint8_t* w = kmers + q * (uint64_t)k;
    uint32_t c = code_of(w[k - 1]);
    uint64_t lo = e.h.pc[c];
    uint64_t hi = lo + (e.emulate_tail ? e.h.init_occ[c] : e.h.tot[c]) - 1;
    for (int j = k - 2; j >= 0; --j) {
      c = code_of(w[j]);
      uint64_t r1, r2;
      occ2(e, c, lo - 1, hi, &r1, &r2); /* as rs_grp_occ2 on the device */
      lo = e.h.pc[c] + r1;
      hi = e.h.pc[c] + r2 - 1;
      if (lo > hi) break;
    }
    lower[q] = lo;
    upper[q] = hi;
  }
}

void emul_prefix_walk_batch(void* p, const uint64_t* rows, uint64_t n, uint32_t max_steps, uint64_t* lexrank,
                            uint32_t* steps) {
  const Emul& e = *(Emul*)p;
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t idx = rows[i];
    uint32_t s = 0;
    while (true) {
      uint64_t nxt;
      const uint32_t sym = lf_step(e, idx, &nxt);
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

}  // extern "C"
