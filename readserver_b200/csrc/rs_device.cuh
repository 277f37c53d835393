/* Lane-group rank / access / select over 128-byte blocks on the device.
 *
 * A query (or walker) is served by LANES in {1, 2, 4, 8} adjacent lanes of a warp; each lane loads
 * 128/LANES bytes of a block with 128-bit loads and decodes its own words (rs_swar.h); with more
 * than one lane, a shuffle scan of the lanes' symbol totals places each lane inside the block and a
 * shuffle reduction adds the partial counts.  LANES = 1 needs no shuffles at all and executes the
 * fewest instructions per rank; LANES = 8 makes one fully coalesced 128-byte request per block.
 *
 * Replaces RLEBWT::getOcc / getChar / getOccAt (src/bwt/rlebwt.cpp:202-301) and the BPTree descent
 * behind them (include/bwt/BPTree.h:50-187).  All arguments of the rs_grp_* functions must be uniform
 * across the lanes of a group; different groups of a warp may diverge (shuffles use the group mask).
 */
#ifndef RS_DEVICE_CUH
#define RS_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"
#include "rs_swar.h"

#define RS_HDR_WORDS (RS_HDR_BYTES / 4) /* 10 */

template <int LANES>
struct rs_grp {
  static constexpr int CH = 8 / LANES;   /* 16-byte chunks per lane */
  static constexpr int NW = 32 / LANES;  /* 32-bit words per lane   */
  unsigned mask;                         /* lanes of this group     */
  int g;                                 /* lane position in group  */
  __device__ __forceinline__ rs_grp() {
    const int lane = threadIdx.x & 31;
    g = lane & (LANES - 1);
    mask = ((LANES == 32) ? 0xFFFFFFFFu : ((1u << LANES) - 1u)) << (lane & ~(LANES - 1));
  }
};

/* A block held by a group: this lane's raw chunks and its payload words (header words zeroed). */
template <int LANES>
struct rs_blk {
  uint4 v[8 / LANES];
  uint32_t w[32 / LANES];
};

template <int LANES>
__device__ __forceinline__ void rs_blk_load(const rs_index_dev& ix, uint64_t b, const rs_grp<LANES>& G, rs_blk<LANES>* B) {
  const uint4* p = reinterpret_cast<const uint4*>(ix.blocks) + b * 8 + G.g * rs_grp<LANES>::CH;
#pragma unroll
  for (int i = 0; i < rs_grp<LANES>::CH; ++i) B->v[i] = __ldg(p + i);
}

/* payload words of the lane: word wi of the block is payload iff wi >= 10 */
template <int LANES>
__device__ __forceinline__ void rs_blk_words(const rs_grp<LANES>& G, rs_blk<LANES>* B) {
  constexpr int NW = rs_grp<LANES>::NW;
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const uint4& q = B->v[j >> 2];
    const uint32_t x = (j & 3) == 0 ? q.x : (j & 3) == 1 ? q.y : (j & 3) == 2 ? q.z : q.w;
    if (LANES == 1) {
      B->w[j] = j >= RS_HDR_WORDS ? x : 0u;
    } else {
      /* lane g, word j is block word g*NW + j: payload iff g*NW + j >= 10 */
      const int first_payload_lane = j >= RS_HDR_WORDS ? 0 : (RS_HDR_WORDS - j + NW - 1) / NW;
      B->w[j] = (G.g >= first_payload_lane) ? x : 0u;
    }
  }
}

template <int LANES>
__device__ __forceinline__ uint32_t rs_grp_exscan(const rs_grp<LANES>& G, uint32_t v, uint32_t* total) {
  if (LANES == 1) {
    *total = v;
    return 0;
  }
  uint32_t s = v;
#pragma unroll
  for (int d = 1; d < LANES; d <<= 1) {
    const uint32_t t = __shfl_up_sync(G.mask, s, d, LANES);
    if (G.g >= d) s += t;
  }
  *total = __shfl_sync(G.mask, s, LANES - 1, LANES);
  return s - v;
}

template <int LANES>
__device__ __forceinline__ uint32_t rs_grp_sum(const rs_grp<LANES>& G, uint32_t v) {
#pragma unroll
  for (int d = LANES / 2; d > 0; d >>= 1) v += __shfl_xor_sync(G.mask, v, d, LANES);
  return v;
}

/* header u64 field f (0 start, 1..4 = A,C,G,T counts) of the block; f uniform in the group */
template <int LANES>
__device__ __forceinline__ uint64_t rs_blk_field(const rs_grp<LANES>& G, const rs_blk<LANES>& B, uint32_t f) {
  /* field f = block words 2f, 2f+1 = chunk f>>1, halves (x,y) or (z,w) */
  uint32_t lo, hi;
  if (LANES == 1) {
    const uint4 a = B.v[0], b = B.v[1], c = B.v[2];
    lo = f == 0 ? a.x : f == 1 ? a.z : f == 2 ? b.x : f == 3 ? b.z : c.x;
    hi = f == 0 ? a.y : f == 1 ? a.w : f == 2 ? b.y : f == 3 ? b.w : c.y;
    return ((uint64_t)hi << 32) | lo;
  }
  constexpr int CH = rs_grp<LANES>::CH;
  const uint32_t chunk = f >> 1;           /* 0..2 */
  const int owner = (int)(chunk / CH);     /* lane holding that chunk */
  const uint32_t ci = chunk % CH;
  uint4 q = B.v[0];
#pragma unroll
  for (int i = 1; i < CH; ++i)
    if (ci == (uint32_t)i) q = B.v[i];
  lo = (f & 1) ? q.z : q.x;
  hi = (f & 1) ? q.w : q.y;
  lo = __shfl_sync(G.mask, lo, owner, LANES);
  hi = __shfl_sync(G.mask, hi, owner, LANES);
  return ((uint64_t)hi << 32) | lo;
}

/* occurrences of c (0..4) before the block */
template <int LANES>
__device__ __forceinline__ uint64_t rs_blk_count_before(const rs_grp<LANES>& G, const rs_blk<LANES>& B, uint32_t c,
                                                        uint64_t start) {
  if (c != 0) return rs_blk_field(G, B, c);
  return start - (rs_blk_field(G, B, 1) + rs_blk_field(G, B, 2) + rs_blk_field(G, B, 3) + rs_blk_field(G, B, 4));
}

/* Positions at or past tail_from answer as clamp - 1 (the reference's tail, rs_layout.h). */
__device__ __forceinline__ uint64_t rs_ref_pos(const rs_index_dev& ix, uint64_t i) {
  return i >= ix.tail_from ? ix.clamp_c - 1 : i;
}
__device__ __forceinline__ uint64_t rs_ref_pos_char(const rs_index_dev& ix, uint64_t i) {
  return i >= ix.tail_from_char ? ix.clamp_char - 1 : i;
}

/* Directory lookup: block holding symbol position pos (< n).  *more: the slot has more boundaries
 * than the entry records and pos lies past the third, so the block must be verified. */
__device__ __forceinline__ uint64_t rs_dir_block(uint64_t e, uint32_t o, bool* more) {
  const uint32_t hi = (uint32_t)(e >> 32);
  const uint32_t b1 = hi & RS_SLOT_MASK, b2 = (hi >> 10) & RS_SLOT_MASK, b3 = (hi >> 20) & RS_SLOT_MASK;
  const uint32_t add = (uint32_t)(b1 != 0 && o >= b1) + (uint32_t)(b2 != 0 && o >= b2) + (uint32_t)(b3 != 0 && o >= b3);
  *more = (hi >> 30) != 0 && add == 3;
  return (uint64_t)(uint32_t)e + add;
}

/* A block settled for one position: decoded words plus the group-wide facts. */
template <int LANES>
struct rs_hit {
  rs_blk<LANES> B;
  uint32_t lane_base; /* block symbols before this lane's words               */
  uint32_t off;       /* 1-based rank of the position inside the block        */
  uint64_t start;     /* symbols before the block                              */
};

/* Decode the block already loaded into H->B for position pos; walk forward while pos lies past its
 * end (only possible when the directory entry flagged `more`). */
template <int LANES>
__device__ __forceinline__ void rs_grp_settle(const rs_index_dev& ix, const rs_grp<LANES>& G, uint64_t pos, uint64_t b,
                                              bool verify, rs_hit<LANES>* H) {
  while (true) {
    rs_blk_words(G, &H->B);
    H->start = rs_blk_field(G, H->B, 0);
    const uint64_t off = pos - H->start + 1;
    uint32_t span = 0;
    if (LANES > 1 || verify) H->lane_base = rs_grp_exscan(G, rs_words_total(H->B.w), &span);
    else H->lane_base = 0;
    if (verify && off > span) { /* group-uniform */
      ++b;
      rs_blk_load(ix, b, G, &H->B);
      continue;
    }
    H->off = (uint32_t)off;
    return;
  }
}

template <int LANES>
__device__ __forceinline__ uint64_t rs_hit_rank(const rs_grp<LANES>& G, const rs_hit<LANES>& H, uint32_t c) {
  uint32_t part, unused;
  rs_words_rank<rs_grp<LANES>::NW, false>(H.B.w, c, (int32_t)H.off - (int32_t)H.lane_base, 0, &part, &unused);
  return rs_blk_count_before(G, H.B, c, H.start) + rs_grp_sum(G, part);
}

/* symbol code at the settled position */
template <int LANES>
__device__ __forceinline__ uint32_t rs_hit_symbol(const rs_grp<LANES>& G, const rs_hit<LANES>& H) {
  const uint32_t mine = rs_words_symbol(H.B.w, (int32_t)H.off - (int32_t)H.lane_base);
  uint32_t sym = mine;
  if (LANES > 1) {
    const unsigned owners = __ballot_sync(G.mask, mine != 0xFFu) & G.mask;
    sym = __shfl_sync(G.mask, mine, (__ffs(owners) - 1) & (LANES - 1), LANES);
  }
  return sym > 4u ? 0u : sym; /* > 4 is unreachable for a well-formed index; keeps pc[] in bounds */
}

template <int LANES>
__device__ __forceinline__ void rs_grp_fetch(const rs_index_dev& ix, const rs_grp<LANES>& G, uint64_t pos, rs_hit<LANES>* H) {
  bool more;
  const uint64_t b = rs_dir_block(__ldg(ix.dir + (pos >> RS_SLOT_SHIFT)), (uint32_t)pos & RS_SLOT_MASK, &more);
  rs_blk_load(ix, b, G, &H->B);
  rs_grp_settle(ix, G, pos, b, more, H);
}

/* getOcc(c, i) with the reference's semantics (i = 2^64-1 -> 0; tail clamp), rlebwt.cpp:268-301. */
template <int LANES>
__device__ __forceinline__ uint64_t rs_grp_occ(const rs_index_dev& ix, const rs_grp<LANES>& G, uint32_t c, uint64_t i) {
  const bool none = (i == ~0ull);
  uint64_t pos = none ? 0 : rs_ref_pos(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  rs_hit<LANES> H;
  rs_grp_fetch(ix, G, pos, &H);
  const uint64_t r = rs_hit_rank(G, H, c);
  return none ? 0 : r;
}

/* getChar(i) with the reference's semantics, rlebwt.cpp:202-227. */
template <int LANES>
__device__ __forceinline__ uint32_t rs_grp_char(const rs_index_dev& ix, const rs_grp<LANES>& G, uint64_t i) {
  uint64_t pos = rs_ref_pos_char(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  rs_hit<LANES> H;
  rs_grp_fetch(ix, G, pos, &H);
  return rs_hit_symbol(G, H);
}

/* One backward-search step (updateInterval, src/bwt/query.cpp:11-15): getOcc(c, i1) and getOcc(c, i2),
 * i1 <= i2 or i1 = 2^64-1.  Both directory entries and both blocks are requested before anything is
 * decoded; when the two positions share a block it is loaded and scanned once for both. */
template <int LANES>
__device__ __forceinline__ void rs_grp_occ2(const rs_index_dev& ix, const rs_grp<LANES>& G, uint32_t c, uint64_t i1,
                                            uint64_t i2, uint64_t* r1, uint64_t* r2) {
  const bool none1 = (i1 == ~0ull), none2 = (i2 == ~0ull);
  uint64_t p1 = none1 ? 0 : rs_ref_pos(ix, i1);
  uint64_t p2 = none2 ? 0 : rs_ref_pos(ix, i2);
  if (p1 >= ix.n) p1 = ix.n - 1;
  if (p2 >= ix.n) p2 = ix.n - 1;
  const uint64_t e1 = __ldg(ix.dir + (p1 >> RS_SLOT_SHIFT));
  const uint64_t e2 = __ldg(ix.dir + (p2 >> RS_SLOT_SHIFT));
  bool m1, m2;
  const uint64_t b1 = rs_dir_block(e1, (uint32_t)p1 & RS_SLOT_MASK, &m1);
  const uint64_t b2 = rs_dir_block(e2, (uint32_t)p2 & RS_SLOT_MASK, &m2);
  rs_hit<LANES> H;
  rs_blk<LANES> B2;
  const bool split = (b1 != b2) || m1 || m2 || p2 < p1; /* group-uniform */
  rs_blk_load(ix, b1, G, &H.B);
  if (split) rs_blk_load(ix, b2, G, &B2);
  /* block of p1: one scan answers p1 and, when it lives in the same block, p2 */
  rs_grp_settle(ix, G, p1, b1, m1, &H);
  const int32_t t1 = (int32_t)H.off - (int32_t)H.lane_base;
  const int32_t t2 = split ? -1 : t1 + (int32_t)(uint32_t)(p2 - p1);
  uint32_t a1, a2;
  rs_words_rank<rs_grp<LANES>::NW, true>(H.B.w, c, t1, t2, &a1, &a2);
  const uint64_t before = rs_blk_count_before(G, H.B, c, H.start);
  if (LANES > 1) {
    a1 = rs_grp_sum(G, a1);
    a2 = rs_grp_sum(G, a2);
  }
  uint64_t x2 = before + a2;
  if (split) { /* group-uniform: p2 sits in another block */
    H.B = B2;
    rs_grp_settle(ix, G, p2, b2, m2, &H);
    x2 = rs_hit_rank(G, H, c);
  }
  *r1 = none1 ? 0 : before + a1;
  *r2 = none2 ? 0 : x2;
}

/* One block scan for up to two targets: rank of c at position pos and, when rel2 >= 0, at pos + rel2
 * (which the caller knows to lie in the same block).  Building block of K1's one-scan-per-trip loop. */
template <int LANES>
__device__ __forceinline__ void rs_grp_scan2(const rs_index_dev& ix, const rs_grp<LANES>& G, uint32_t c, uint64_t pos,
                                             uint64_t b, bool verify, int32_t rel2, uint64_t* x1, uint64_t* x2) {
  rs_hit<LANES> H;
  rs_blk_load(ix, b, G, &H.B);
  rs_grp_settle(ix, G, pos, b, verify, &H);
  const int32_t t1 = (int32_t)H.off - (int32_t)H.lane_base;
  const int32_t t2 = rel2 < 0 ? -1 : t1 + rel2;
  uint32_t a1, a2;
  rs_words_rank<rs_grp<LANES>::NW, true>(H.B.w, c, t1, t2, &a1, &a2);
  const uint64_t before = rs_blk_count_before(G, H.B, c, H.start);
  if (LANES > 1) {
    a1 = rs_grp_sum(G, a1);
    a2 = rs_grp_sum(G, a2);
  }
  *x1 = before + a1;
  *x2 = before + a2;
}

/* One LF step of extractPrefix (src/bwt/query.cpp:49-51): b = getChar(idx); idx' = PC(b) + getOcc(b, idx-1).
 * Returns the symbol code (0 = '$') and writes idx'.  Bit-exact with the reference including its tail. */
template <int LANES>
__device__ __forceinline__ uint32_t rs_grp_lf_step(const rs_index_dev& ix, const rs_grp<LANES>& G, uint64_t idx,
                                                   uint64_t* next) {
  uint64_t pc = rs_ref_pos_char(ix, idx);
  if (pc >= ix.n) pc = ix.n - 1;
  rs_hit<LANES> H;
  rs_grp_fetch(ix, G, pc, &H);
  const uint32_t sym = rs_hit_symbol(G, H);
  /* getOcc(sym, idx-1): when idx-1 resolves to the position just before pc it is rank(sym, pc) - 1 */
  const bool none = (idx == 0);
  const uint64_t pp = none ? 0 : rs_ref_pos(ix, idx - 1);
  uint64_t occ;
  if (!none && pp + 1 == pc) {
    occ = rs_hit_rank(G, H, sym) - 1;
  } else {
    occ = rs_grp_occ(ix, G, sym, idx - 1); /* reference tail region, or idx == 0 */
  }
  *next = ix.pc[sym] + occ;
  return sym;
}

/* occurrences of c before block b, read straight from its header (no decode) */
__device__ __forceinline__ uint64_t rs_hdr_count_before(const rs_index_dev& ix, uint64_t b, uint32_t c) {
  const uint64_t* h = reinterpret_cast<const uint64_t*>(ix.blocks) + b * (RS_BLOCK_BYTES / 8);
  if (c != 0) return __ldg(h + c);
  return __ldg(h) - (__ldg(h + 1) + __ldg(h + 2) + __ldg(h + 3) + __ldg(h + 4));
}

/* getOccAt(c, r) (rlebwt.cpp:233-266): position of the r-th occurrence of c, 1 <= r <= tot[c].
 * The reference's select never consults m_vSum, so there is no tail behaviour to reproduce
 * (SURVEY.md A8).  r outside [1, tot[c]] yields 2^64-1. */
template <int LANES>
__device__ __forceinline__ uint64_t rs_grp_occ_at(const rs_index_dev& ix, const rs_grp<LANES>& G, uint32_t c, uint64_t r) {
  const bool bad = (r == 0 || r > ix.tot[c]);
  if (bad) r = 1;
  if (ix.tot[c] == 0) return ~0ull; /* uniform: symbol absent from the BWT */
  uint64_t b = __ldg(ix.sel[c] + ((r - 1) >> ix.sel_shift[c]));
  /* skip blocks that end before the r-th occurrence: their successor's header says so */
  while (b + 1 < ix.nblocks && rs_hdr_count_before(ix, b + 1, c) < r) ++b;
  rs_hit<LANES> H;
  rs_blk_load(ix, b, G, &H.B);
  rs_blk_words(G, &H.B);
  H.start = rs_blk_field(G, H.B, 0);
  const uint64_t before = rs_blk_count_before(G, H.B, c, H.start);
  const uint32_t want = (uint32_t)(r - before);
  uint32_t span, ctot;
  const uint32_t lane_base = rs_grp_exscan(G, rs_words_total(H.B.w), &span);
  const uint32_t cbase = rs_grp_exscan(G, rs_words_count_all(H.B.w, c), &ctot);
  const uint32_t p = rs_words_select(H.B.w, c, (int32_t)want - (int32_t)cbase);
  uint32_t in_block = p ? lane_base + p : 0u; /* 1-based position inside the block, from the owning lane */
  if (LANES > 1) {
    const unsigned owners = __ballot_sync(G.mask, p != 0) & G.mask;
    in_block = __shfl_sync(G.mask, in_block, (__ffs(owners) - 1) & (LANES - 1), LANES);
  }
  return bad ? ~0ull : H.start + in_block - 1;
}

/* getF(i) (rlebwt.cpp:307-314) as a symbol code */
__device__ __forceinline__ uint32_t rs_f_code(const rs_index_dev& ix, uint64_t i) {
  uint32_t ci = 0;
#pragma unroll
  for (int c = 1; c < RS_NSYM; ++c) ci += (ix.pc[c] <= i) ? 1u : 0u;
  return ci;
}

/* One step of extractPostfix (src/bwt/query.cpp:69-75): f = getF(idx); idx' = getOccAt(f, idx - PC(f) + 1). */
template <int LANES>
__device__ __forceinline__ uint32_t rs_grp_fl_step(const rs_index_dev& ix, const rs_grp<LANES>& G, uint64_t idx,
                                                   uint64_t* next) {
  const uint32_t f = rs_f_code(ix, idx);
  *next = rs_grp_occ_at(ix, G, f, idx - ix.pc[f] + 1);
  return f;
}

#endif
