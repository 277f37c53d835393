/* Lane-group (8 lanes x 16 B = one 128-byte block) rank / access / select on the device.
 *
 * Replaces RLEBWT::getOcc / getChar / getOccAt (src/bwt/rlebwt.cpp:202-301) and the BPTree descent
 * behind them (include/bwt/BPTree.h:50-187).  All arguments of the rs_grp_* functions must be uniform
 * across the 8 lanes of a group; different groups of a warp may diverge (shuffles use the group mask).
 */
#ifndef RS_DEVICE_CUH
#define RS_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"
#include "rs_swar.h"

struct rs_grp {
  unsigned mask; /* the 8 lanes of this group */
  int g;         /* lane position in the group */
  int base;      /* first lane of the group */
};

__device__ __forceinline__ rs_grp rs_grp_init() {
  rs_grp G;
  const int lane = threadIdx.x & 31;
  G.g = lane & (RS_GROUP - 1);
  G.base = lane & ~(RS_GROUP - 1);
  G.mask = 0xFFu << G.base;
  return G;
}

__device__ __forceinline__ uint4 rs_ld_block(const rs_index_dev& ix, uint64_t b, int g) {
  return __ldg(reinterpret_cast<const uint4*>(ix.blocks) + b * RS_GROUP + g);
}

/* Reference tail behaviour: positions in the last m_vSum window at or past the clamp
 * answer as clamp - 1 (clamp_c for getOcc, clamp_char for getChar; see rs_layout.h). */
__device__ __forceinline__ uint64_t rs_ref_pos(const rs_index_dev& ix, uint64_t i) {
  if (ix.emulate_tail && ((i + 1) >> 16) >= ix.vlast && i >= ix.clamp_c) i = ix.clamp_c - 1;
  return i;
}
__device__ __forceinline__ uint64_t rs_ref_pos_char(const rs_index_dev& ix, uint64_t i) {
  if (ix.emulate_tail && ((i + 1) >> 16) >= ix.vlast && i >= ix.clamp_char) i = ix.clamp_char - 1;
  return i;
}

/* Directory lookup: block holding symbol position pos (< n).  *more: the slot has more boundaries
 * than the entry records and pos lies past the third, so the block must be verified. */
__device__ __forceinline__ uint64_t rs_dir_block(uint64_t e, uint32_t o, bool* more) {
  const uint32_t b1 = (uint32_t)(e >> 32) & RS_SLOT_MASK;
  const uint32_t b2 = (uint32_t)(e >> 42) & RS_SLOT_MASK;
  const uint32_t b3 = (uint32_t)(e >> 52) & RS_SLOT_MASK;
  const uint32_t add = (uint32_t)(b1 != 0 && o >= b1) + (uint32_t)(b2 != 0 && o >= b2) + (uint32_t)(b3 != 0 && o >= b3);
  *more = (e & RS_DIR_MORE) != 0 && add == 3;
  return (uint64_t)(uint32_t)e + add;
}

/* Inclusive scan of `v` across the group; returns the exclusive prefix, *total = group sum. */
__device__ __forceinline__ uint32_t rs_grp_exscan(const rs_grp& G, uint32_t v, uint32_t* total) {
  uint32_t s = v;
#pragma unroll
  for (int d = 1; d < RS_GROUP; d <<= 1) {
    const uint32_t t = __shfl_up_sync(G.mask, s, d, RS_GROUP);
    if (G.g >= d) s += t;
  }
  *total = __shfl_sync(G.mask, s, RS_GROUP - 1, RS_GROUP);
  return s - v;
}

__device__ __forceinline__ uint32_t rs_grp_sum(const rs_grp& G, uint32_t v) {
#pragma unroll
  for (int d = RS_GROUP / 2; d > 0; d >>= 1) v += __shfl_xor_sync(G.mask, v, d, RS_GROUP);
  return v;
}

__device__ __forceinline__ uint64_t rs_grp_bcast64(const rs_grp& G, uint32_t lo, uint32_t hi, int src) {
  lo = __shfl_sync(G.mask, lo, src, RS_GROUP);
  hi = __shfl_sync(G.mask, hi, src, RS_GROUP);
  return ((uint64_t)hi << 32) | lo;
}

/* Header fields of the block whose 16-byte pieces the lanes hold in v. */
__device__ __forceinline__ uint64_t rs_hdr_start(const rs_grp& G, const uint4& v) {
  return rs_grp_bcast64(G, v.x, v.y, 0);
}

/* occurrences of c (0..4) before the block */
__device__ __forceinline__ uint64_t rs_hdr_count(const rs_grp& G, const uint4& v, uint32_t c, uint64_t start) {
  if (c != 0) {
    const uint32_t f = c - 1; /* A,C in lanes 0(z,w) .. : field f+1 of the header */
    /* header u64 fields: 0 start | 1 A | 2 C | 3 G | 4 T ; two fields per lane */
    const uint32_t field = f + 1;
    const bool hi_half = field & 1;
    const uint32_t lo = hi_half ? v.z : v.x, hi = hi_half ? v.w : v.y;
    return rs_grp_bcast64(G, lo, hi, field >> 1);
  }
  const uint64_t a = rs_grp_bcast64(G, v.z, v.w, 0);
  const uint64_t cc = rs_grp_bcast64(G, v.x, v.y, 1);
  const uint64_t gg = rs_grp_bcast64(G, v.z, v.w, 1);
  const uint64_t t = rs_grp_bcast64(G, v.x, v.y, 2);
  return start - (a + cc + gg + t);
}

/* A block resolved for one symbol position: the lane's decoded bytes plus the group-wide facts. */
struct rs_hit {
  uint4 v;            /* this lane's 16 bytes of the block                     */
  rs_lane L;
  uint32_t lane_base; /* block symbols before this lane's bytes               */
  uint32_t off;       /* 1-based rank of the position inside the block        */
  uint64_t start;     /* symbols before the block                              */
};

/* Decode the block in H.v for position pos, walking forward while pos lies past its end (only
 * possible when the directory entry flagged `more`). */
__device__ __forceinline__ void rs_grp_settle(const rs_index_dev& ix, const rs_grp& G, uint64_t pos, uint64_t b,
                                              bool verify, rs_hit* H) {
  while (true) {
    rs_lane_prepare(H->v.x, H->v.y, H->v.z, H->v.w, G.g, &H->L);
    uint32_t span;
    H->lane_base = rs_grp_exscan(G, H->L.total, &span);
    H->start = rs_hdr_start(G, H->v);
    const uint64_t off = pos - H->start + 1;
    if (verify && off > span) { /* group-uniform */
      ++b;
      H->v = rs_ld_block(ix, b, G.g);
      continue;
    }
    H->off = (uint32_t)off;
    return;
  }
}

/* occurrences of c in BWT[0..pos] given the settled block of pos */
__device__ __forceinline__ uint64_t rs_hit_rank(const rs_grp& G, const rs_hit& H, uint32_t c) {
  const uint32_t part = rs_lane_count(&H.L, c, (int32_t)H.off - (int32_t)H.lane_base);
  return rs_hdr_count(G, H.v, c, H.start) + rs_grp_sum(G, part);
}

/* symbol code at pos given the settled block of pos */
__device__ __forceinline__ uint32_t rs_hit_symbol(const rs_grp& G, const rs_hit& H) {
  const uint32_t mine = rs_lane_symbol(&H.L, (int32_t)H.off - (int32_t)H.lane_base);
  const unsigned owners = __ballot_sync(G.mask, mine != 0xFFu) & G.mask;
  const uint32_t sym = __shfl_sync(G.mask, mine, (__ffs(owners) - 1) & (RS_GROUP - 1), RS_GROUP);
  return sym > 4u ? 0u : sym; /* > 4 is unreachable for a well-formed index; keeps pc[] in bounds */
}

__device__ __forceinline__ void rs_grp_fetch(const rs_index_dev& ix, const rs_grp& G, uint64_t pos, rs_hit* H) {
  bool more;
  const uint64_t b = rs_dir_block(__ldg(ix.dir + (pos >> RS_SLOT_SHIFT)), (uint32_t)pos & RS_SLOT_MASK, &more);
  H->v = rs_ld_block(ix, b, G.g);
  rs_grp_settle(ix, G, pos, b, more, H);
}

/* getOcc(c, i) with the reference's semantics (i = 2^64-1 -> 0; tail clamp), rlebwt.cpp:268-301. */
__device__ __forceinline__ uint64_t rs_grp_occ(const rs_index_dev& ix, const rs_grp& G, uint32_t c, uint64_t i) {
  const bool none = (i == ~0ull);
  uint64_t pos = none ? 0 : rs_ref_pos(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  rs_hit H;
  rs_grp_fetch(ix, G, pos, &H);
  const uint64_t r = rs_hit_rank(G, H, c);
  return none ? 0 : r;
}

/* getChar(i) with the reference's semantics, rlebwt.cpp:202-227. */
__device__ __forceinline__ uint32_t rs_grp_char(const rs_index_dev& ix, const rs_grp& G, uint64_t i) {
  uint64_t pos = rs_ref_pos_char(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  rs_hit H;
  rs_grp_fetch(ix, G, pos, &H);
  return rs_hit_symbol(G, H);
}

/* Two ranks of the same symbol with both loads in flight together (one backward-search step:
 * updateInterval, src/bwt/query.cpp:11-15). */
__device__ __forceinline__ void rs_grp_occ2(const rs_index_dev& ix, const rs_grp& G, uint32_t c, uint64_t i1,
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
  rs_hit H1, H2;
  H1.v = rs_ld_block(ix, b1, G.g);
  H2.v = rs_ld_block(ix, b2, G.g);
  rs_grp_settle(ix, G, p1, b1, m1, &H1);
  const uint64_t a1 = rs_hit_rank(G, H1, c);
  rs_grp_settle(ix, G, p2, b2, m2, &H2);
  const uint64_t a2 = rs_hit_rank(G, H2, c);
  *r1 = none1 ? 0 : a1;
  *r2 = none2 ? 0 : a2;
}

/* One LF step of extractPrefix (src/bwt/query.cpp:49-51): b = getChar(idx); idx' = PC(b) + getOcc(b, idx-1).
 * Returns the symbol code (0 = '$') and writes idx'.  Bit-exact with the reference including its tail. */
__device__ __forceinline__ uint32_t rs_grp_lf_step(const rs_index_dev& ix, const rs_grp& G, uint64_t idx, uint64_t* next) {
  uint64_t pc = rs_ref_pos_char(ix, idx);
  if (pc >= ix.n) pc = ix.n - 1;
  rs_hit H;
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

#endif
