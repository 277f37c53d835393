/* Rank / access / select over the 128-byte blocks of rs_layout.h, one thread per query or walker.
 *
 * Every operation resolves its block through the position (or select) directory, reads the 32-byte
 * header sector, picks the sub-block from the header's bounds and reads that one 32-byte sector:
 * four 128-bit loads and a 6-word SWAR scan (rs_swar.h) per rank.  The header load carries an
 * L2::128B prefetch hint, so that the sub-block sector is an L2 hit when the line came from DRAM.
 *
 * The same source compiles for the sm_100a kernels and for the host: tests/emul replays it on host
 * memory (pointers of rs_index_dev into the host builder's arrays) against the oracle at every
 * position.  The product library never runs it on the host.
 *
 * Replaces RLEBWT::getOcc / getChar / getOccAt / getF (src/bwt/rlebwt.cpp:202-314) and the BPTree
 * descent behind them (include/bwt/BPTree.h:50-187).
 */
#ifndef RS_OPS_H
#define RS_OPS_H

#include <stdint.h>
#include <string.h>

#include "rs_layout.h"
#include "rs_swar.h"

struct rs_u4 {
  uint32_t x, y, z, w;
};

/* 16-byte load of index data (read-only for the lifetime of the index) */
RS_HD rs_u4 rs_ld128(const void* p) {
#if defined(__CUDA_ARCH__)
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  return rs_u4{v.x, v.y, v.z, v.w};
#else
  rs_u4 v;
  memcpy(&v, p, 16);
  return v;
#endif
}

/* the same with a hint to bring the whole 128-byte line into L2 on a miss */
RS_HD rs_u4 rs_ld128_line(const void* p) {
#if defined(__CUDA_ARCH__) && !defined(RS_NO_LINE_PREFETCH)
  rs_u4 v;
  asm volatile("ld.global.nc.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(__cvta_generic_to_global(p)));
  return v;
#else
  return rs_ld128(p);
#endif
}

RS_HD uint64_t rs_ld64(const uint64_t* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

RS_HD uint32_t rs_ld32(const uint32_t* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

RS_HD const uint8_t* rs_block_ptr(const rs_index_dev& ix, uint64_t b) {
  return reinterpret_cast<const uint8_t*>(ix.blocks) + b * RS_BLOCK_BYTES;
}

/* Positions at or past tail_from answer as clamp - 1 (the reference's tail, rs_layout.h). */
RS_HD uint64_t rs_ref_pos(const rs_index_dev& ix, uint64_t i) { return i >= ix.tail_from ? ix.clamp_c - 1 : i; }
RS_HD uint64_t rs_ref_pos_char(const rs_index_dev& ix, uint64_t i) {
  return i >= ix.tail_from_char ? ix.clamp_char - 1 : i;
}

/* Block holding symbol position pos (< n). */
RS_HD uint64_t rs_find_block(const rs_index_dev& ix, uint64_t pos) {
  const uint64_t e = rs_ld64(ix.dir + (pos >> RS_SLOT_SHIFT));
  const uint32_t o = (uint32_t)pos & RS_SLOT_MASK;
  const uint32_t hi = (uint32_t)(e >> 32);
  const uint32_t b1 = hi & RS_SLOT_MASK, b2 = (hi >> 10) & RS_SLOT_MASK, b3 = (hi >> 20) & RS_SLOT_MASK;
  const uint32_t add = (uint32_t)(b1 != 0 && o >= b1) + (uint32_t)(b2 != 0 && o >= b2) + (uint32_t)(b3 != 0 && o >= b3);
  uint64_t b = (uint64_t)(uint32_t)e + add;
  if ((hi >> 30) != 0 && add == 3) {
    /* more than three boundaries in the slot and pos lies past the third: compare start fields */
    while (b + 1 < ix.nblocks && rs_ld64(reinterpret_cast<const uint64_t*>(rs_block_ptr(ix, b + 1))) <= pos) ++b;
  }
  return b;
}

/* header sector of a block */
struct rs_hdr {
  uint64_t start;
  uint32_t cl[4];
  uint32_t ch;
  uint32_t bounds;
};

RS_HD rs_hdr rs_load_hdr(const rs_index_dev& ix, uint64_t b) {
  const uint8_t* p = rs_block_ptr(ix, b);
  const rs_u4 a = rs_ld128_line(p), c = rs_ld128(p + 16);
  rs_hdr h;
  h.start = ((uint64_t)a.y << 32) | a.x;
  h.cl[0] = a.z;
  h.cl[1] = a.w;
  h.cl[2] = c.x;
  h.cl[3] = c.y;
  h.ch = c.z;
  h.bounds = c.w;
  return h;
}

/* sub-block (0..2) holding the off-th symbol of the block, off 1-based */
RS_HD uint32_t rs_hdr_sub(const rs_hdr& h, uint32_t off) {
  const uint32_t s1 = h.bounds & RS_REC_MASK, s2 = (h.bounds >> RS_REC_BITS) & RS_REC_MASK;
  return (uint32_t)(off > s1) + (uint32_t)(off > s2);
}

/* occurrences of c (0..4) before the block */
RS_HD uint64_t rs_hdr_count(const rs_hdr& h, uint32_t c) {
  if (c != 0) {
    const uint32_t k = c - 1;
    const uint32_t lo = k == 0 ? h.cl[0] : k == 1 ? h.cl[1] : k == 2 ? h.cl[2] : h.cl[3];
    return ((uint64_t)((h.ch >> (8 * k)) & 0xFFu) << 32) | lo;
  }
  const uint64_t lo = (uint64_t)h.cl[0] + h.cl[1] + h.cl[2] + h.cl[3];
  const uint64_t hi = (h.ch & 0xFFu) + ((h.ch >> 8) & 0xFFu) + ((h.ch >> 16) & 0xFFu) + (h.ch >> 24);
  return h.start - (lo + (hi << 32));
}

/* one sub-block sector */
struct rs_sub {
  uint64_t rec;
  uint32_t w[RS_SUB_WORDS];
};

RS_HD rs_sub rs_load_sub(const rs_index_dev& ix, uint64_t b, uint32_t j) {
  const uint8_t* p = rs_block_ptr(ix, b) + RS_HDR_BYTES + j * 32u;
  const rs_u4 a = rs_ld128(p), c = rs_ld128(p + 16);
  rs_sub s;
  s.rec = ((uint64_t)a.y << 32) | a.x;
  s.w[0] = a.z;
  s.w[1] = a.w;
  s.w[2] = c.x;
  s.w[3] = c.y;
  s.w[4] = c.z;
  s.w[5] = c.w;
  return s;
}

RS_HD uint32_t rs_rec_sym(uint64_t rec) { return (uint32_t)(rec >> RS_REC_SYM_SHIFT) & 0xFFFu; }

/* occurrences of c inside the block before the sub-block whose record is rec */
RS_HD uint32_t rs_rec_count(uint64_t rec, uint32_t c) {
  if (c != 0) return (uint32_t)(rec >> (RS_REC_BITS * (c - 1))) & RS_REC_MASK;
  const uint32_t a = (uint32_t)rec & RS_REC_MASK, cc = (uint32_t)(rec >> 11) & RS_REC_MASK;
  const uint32_t g = (uint32_t)(rec >> 22) & RS_REC_MASK, t = (uint32_t)(rec >> 33) & RS_REC_MASK;
  return rs_rec_sym(rec) - (a + cc + g + t);
}

/* True rank: occurrences of c in BWT[0..pos], pos < n. */
RS_HD uint64_t rs_rank(const rs_index_dev& ix, uint32_t c, uint64_t pos) {
  const uint64_t b = rs_find_block(ix, pos);
  const rs_hdr h = rs_load_hdr(ix, b);
  const uint32_t off = (uint32_t)(pos - h.start) + 1u;
  const rs_sub s = rs_load_sub(ix, b, rs_hdr_sub(h, off));
  uint32_t a, unused;
  rs_words_rank<RS_SUB_WORDS, false>(s.w, c, (int32_t)(off - rs_rec_sym(s.rec)), 0, &a, &unused);
  return rs_hdr_count(h, c) + rs_rec_count(s.rec, c) + a;
}

/* getOcc(c, i) with the reference's semantics (i = 2^64-1 -> 0; tail clamp), rlebwt.cpp:268-301. */
RS_HD uint64_t rs_occ(const rs_index_dev& ix, uint32_t c, uint64_t i) {
  if (i == ~0ull) return 0;
  uint64_t pos = rs_ref_pos(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  return rs_rank(ix, c, pos);
}

/* One block scan for up to two targets: rank of c at p1 and, when try2 and p2 (>= p1) falls in the
 * same sub-block, at p2 as well (*did2).  Building block of K1's one-scan-per-trip loop. */
RS_HD void rs_scan2(const rs_index_dev& ix, uint32_t c, uint64_t b, uint64_t p1, uint64_t p2, bool try2, uint64_t* x1,
                    uint64_t* x2, bool* did2) {
  const rs_hdr h = rs_load_hdr(ix, b);
  const uint32_t off = (uint32_t)(p1 - h.start) + 1u;
  const uint32_t j = rs_hdr_sub(h, off);
  const uint32_t off2 = off + (uint32_t)(p2 - p1);
  const bool two = try2 && rs_hdr_sub(h, off2) == j;
  const rs_sub s = rs_load_sub(ix, b, j);
  const uint32_t sb = rs_rec_sym(s.rec);
  uint32_t a1, a2;
  rs_words_rank<RS_SUB_WORDS, true>(s.w, c, (int32_t)(off - sb), two ? (int32_t)(off2 - sb) : -1, &a1, &a2);
  const uint64_t before = rs_hdr_count(h, c) + rs_rec_count(s.rec, c);
  *x1 = before + a1;
  *x2 = before + a2;
  *did2 = two;
}

/* updateInterval's two ranks (src/bwt/query.cpp:11-15): getOcc(c, i1) and getOcc(c, i2). */
RS_HD void rs_occ2(const rs_index_dev& ix, uint32_t c, uint64_t i1, uint64_t i2, uint64_t* r1, uint64_t* r2) {
  const bool none1 = (i1 == ~0ull), none2 = (i2 == ~0ull);
  uint64_t p1 = none1 ? 0 : rs_ref_pos(ix, i1);
  uint64_t p2 = none2 ? 0 : rs_ref_pos(ix, i2);
  if (p1 >= ix.n) p1 = ix.n - 1;
  if (p2 >= ix.n) p2 = ix.n - 1;
  const uint64_t b1 = rs_find_block(ix, p1), b2 = rs_find_block(ix, p2);
  uint64_t x1, x2;
  bool did2;
  rs_scan2(ix, c, b1, p1, p2, b1 == b2 && p2 >= p1, &x1, &x2, &did2);
  if (!did2) {
    uint64_t unused;
    bool u2;
    rs_scan2(ix, c, b2, p2, p2, false, &x2, &unused, &u2);
  }
  *r1 = none1 ? 0 : x1;
  *r2 = none2 ? 0 : x2;
}

/* Symbol at pos (< n) and the rank of that symbol at pos, from one scan. */
RS_HD uint32_t rs_symbol_rank(const rs_index_dev& ix, uint64_t pos, uint64_t* rank) {
  const uint64_t b = rs_find_block(ix, pos);
  const rs_hdr h = rs_load_hdr(ix, b);
  const uint32_t off = (uint32_t)(pos - h.start) + 1u;
  const rs_sub s = rs_load_sub(ix, b, rs_hdr_sub(h, off));
  const int32_t t = (int32_t)(off - rs_rec_sym(s.rec));
  uint32_t sym = rs_words_symbol<RS_SUB_WORDS>(s.w, t);
  if (sym > 4u) sym = 0u; /* unreachable for a well-formed index; keeps pc[] in bounds */
  if (rank) {
    uint32_t a, unused;
    rs_words_rank<RS_SUB_WORDS, false>(s.w, sym, t, 0, &a, &unused);
    *rank = rs_hdr_count(h, sym) + rs_rec_count(s.rec, sym) + a;
  }
  return sym;
}

/* getChar(i) with the reference's semantics, rlebwt.cpp:202-227. */
RS_HD uint32_t rs_char(const rs_index_dev& ix, uint64_t i) {
  uint64_t pos = rs_ref_pos_char(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  return rs_symbol_rank(ix, pos, nullptr);
}

/* One LF step of extractPrefix (src/bwt/query.cpp:49-51): b = getChar(idx); idx' = PC(b) + getOcc(b, idx-1).
 * Returns the symbol code (0 = '$') and writes idx'.  Bit-exact with the reference including its tail. */
RS_HD uint32_t rs_lf_step(const rs_index_dev& ix, uint64_t idx, uint64_t* next) {
  uint64_t pc = rs_ref_pos_char(ix, idx);
  if (pc >= ix.n) pc = ix.n - 1;
  uint64_t r;
  const uint32_t sym = rs_symbol_rank(ix, pc, &r);
  /* getOcc(sym, idx-1): when idx-1 resolves to the position just before pc it is rank(sym, pc) - 1 */
  const bool none = (idx == 0);
  const uint64_t pp = none ? 0 : rs_ref_pos(ix, idx - 1);
  uint64_t occ;
  if (!none && pp + 1 == pc) occ = r - 1;
  else occ = rs_occ(ix, sym, idx - 1); /* reference tail region, or idx == 0 */
  *next = ix.pc[sym] + occ;
  return sym;
}

/* occurrences of c before block b, from its header sector */
RS_HD uint64_t rs_block_count_before(const rs_index_dev& ix, uint64_t b, uint32_t c) {
  const uint8_t* p = rs_block_ptr(ix, b);
  if (c != 0) {
    const uint32_t lo = rs_ld32(reinterpret_cast<const uint32_t*>(p + 8) + (c - 1));
    const uint32_t ch = rs_ld32(reinterpret_cast<const uint32_t*>(p + 24));
    return ((uint64_t)((ch >> (8 * (c - 1))) & 0xFFu) << 32) | lo;
  }
  return rs_hdr_count(rs_load_hdr(ix, b), 0);
}

/* getOccAt(c, r) (rlebwt.cpp:233-266): position of the r-th occurrence of c, 1 <= r <= tot[c].
 * The reference's select never consults m_vSum, so there is no tail behaviour to reproduce
 * (SURVEY.md A8).  r outside [1, tot[c]] yields 2^64-1. */
RS_HD uint64_t rs_occ_at(const rs_index_dev& ix, uint32_t c, uint64_t r) {
  if (r == 0 || r > ix.tot[c]) return ~0ull;
  uint64_t b = rs_ld32(ix.sel[c] + ((r - 1) >> ix.sel_shift[c]));
  /* skip blocks that end before the r-th occurrence: their successor's header says so */
  while (b + 1 < ix.nblocks && rs_block_count_before(ix, b + 1, c) < r) ++b;
  const rs_hdr h = rs_load_hdr(ix, b);
  const uint32_t want = (uint32_t)(r - rs_hdr_count(h, c));
  const uint8_t* p = rs_block_ptr(ix, b) + RS_HDR_BYTES;
  const uint32_t c1 = rs_rec_count(rs_ld64(reinterpret_cast<const uint64_t*>(p + 32)), c);
  const uint32_t c2 = rs_rec_count(rs_ld64(reinterpret_cast<const uint64_t*>(p + 64)), c);
  /* empty trailing sub-blocks of the last block repeat the block's totals, so want <= c1 stops early */
  const uint32_t j = (uint32_t)(want > c1) + (uint32_t)(want > c2);
  const rs_sub s = rs_load_sub(ix, b, j);
  const uint32_t q = rs_words_select<RS_SUB_WORDS>(s.w, c, (int32_t)(want - rs_rec_count(s.rec, c)));
  return h.start + rs_rec_sym(s.rec) + q - 1;
}

/* getF(i) (rlebwt.cpp:307-314) as a symbol code */
RS_HD uint32_t rs_f_code(const rs_index_dev& ix, uint64_t i) {
  uint32_t ci = 0;
#pragma unroll
  for (int c = 1; c < RS_NSYM; ++c) ci += (ix.pc[c] <= i) ? 1u : 0u;
  return ci;
}

/* One step of extractPostfix (src/bwt/query.cpp:69-75): f = getF(idx); idx' = getOccAt(f, idx - PC(f) + 1). */
RS_HD uint32_t rs_fl_step(const rs_index_dev& ix, uint64_t idx, uint64_t* next) {
  const uint32_t f = rs_f_code(ix, idx);
  *next = rs_occ_at(ix, f, idx - ix.pc[f] + 1);
  return f;
}

#endif
