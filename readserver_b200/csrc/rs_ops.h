/* Rank / access / select over the 64-byte blocks of rs_layout.h, one thread per query or walker.
 *
 * Every operation resolves its block through the position (or select) directory and reads the whole
 * block with four independent 128-bit loads: header (start + counters) and 9 words of run bytes,
 * scanned with 32-bit SWAR (rs_swar.h).  Two dependent memory accesses and two DRAM sectors per rank.
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

RS_HD uint64_t rs_ld64(const uint64_t* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

/* 8-byte load of a directory word: asks L2 to keep the line (evict-last) -- the position directory is read on
 * every rank and is ~7x smaller than the blocks it competes with for L2.  RS_DIR_KEEP=0 compiles it out. */
#ifndef RS_DIR_KEEP
#define RS_DIR_KEEP 1
#endif
RS_HD uint64_t rs_ld64_keep(const uint64_t* p) {
#if defined(__CUDA_ARCH__) && RS_DIR_KEEP
  uint64_t pol, v;
  asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(__cvta_generic_to_global(p)), "l"(pol));
  return v;
#else
  return rs_ld64(p);
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
  const uint64_t e = rs_ld64_keep(ix.dir + (pos >> RS_SLOT_SHIFT));
  const uint32_t o = (uint32_t)pos & RS_SLOT_MASK;
  const uint32_t hi = (uint32_t)(e >> 32);
  const uint32_t b1 = hi & RS_SLOT_MASK, b2 = (hi >> RS_SLOT_SHIFT) & RS_SLOT_MASK,
                 b3 = (hi >> (2 * RS_SLOT_SHIFT)) & RS_SLOT_MASK;
  const uint32_t add = (uint32_t)(b1 != 0 && o >= b1) + (uint32_t)(b2 != 0 && o >= b2) + (uint32_t)(b3 != 0 && o >= b3);
  uint64_t b = (uint64_t)(uint32_t)e + add;
  if ((hi >> 30) != 0 && add == 3) {
    /* more than three boundaries in the slot and pos lies past the third: compare start fields */
    while (b + 1 < ix.nblocks && rs_ld64(reinterpret_cast<const uint64_t*>(rs_block_ptr(ix, b + 1))) <= pos) ++b;
  }
  return b;
}

/* a whole block: header fields and the 9 payload words */
struct rs_blk {
  uint64_t start;
  uint32_t cl[4];
  uint32_t ch;
  uint32_t w[RS_BLOCK_WORDS];
};

RS_HD rs_blk rs_load_block(const rs_index_dev& ix, uint64_t b) {
  const uint8_t* p = rs_block_ptr(ix, b);
  const rs_u4 v0 = rs_ld128(p), v1 = rs_ld128(p + 16), v2 = rs_ld128(p + 32), v3 = rs_ld128(p + 48);
  rs_blk k;
  k.start = ((uint64_t)v0.y << 32) | v0.x;
  k.cl[0] = v0.z;
  k.cl[1] = v0.w;
  k.cl[2] = v1.x;
  k.cl[3] = v1.y;
  k.ch = v1.z;
  k.w[0] = v1.w;
  k.w[1] = v2.x;
  k.w[2] = v2.y;
  k.w[3] = v2.z;
  k.w[4] = v2.w;
  k.w[5] = v3.x;
  k.w[6] = v3.y;
  k.w[7] = v3.z;
  k.w[8] = v3.w;
  return k;
}

/* occurrences of c (0..4) before the block */
RS_HD uint64_t rs_blk_count(const rs_blk& h, uint32_t c) {
  if (c != 0) {
    const uint32_t k = c - 1;
    const uint32_t lo = k == 0 ? h.cl[0] : k == 1 ? h.cl[1] : k == 2 ? h.cl[2] : h.cl[3];
    return ((uint64_t)((h.ch >> (8 * k)) & 0xFFu) << 32) | lo;
  }
  const uint64_t lo = (uint64_t)h.cl[0] + h.cl[1] + h.cl[2] + h.cl[3];
  const uint64_t hi = (h.ch & 0xFFu) + ((h.ch >> 8) & 0xFFu) + ((h.ch >> 16) & 0xFFu) + (h.ch >> 24);
  return h.start - (lo + (hi << 32));
}

/* One block scan for up to two targets: rank of c at p1 and, when two (p2 >= p1 lies in the same block),
 * at p2 as well.  Building block of K1's one-scan-per-trip loop. */
RS_HD void rs_scan2(const rs_index_dev& ix, uint32_t c, uint64_t b, uint64_t p1, uint64_t p2, bool two, uint64_t* x1,
                    uint64_t* x2) {
  const rs_blk k = rs_load_block(ix, b);
  const uint32_t off = (uint32_t)(p1 - k.start) + 1u;
  uint32_t a1, a2;
  rs_words_rank<RS_BLOCK_WORDS, true>(k.w, c, (int32_t)off, two ? (int32_t)(off + (uint32_t)(p2 - p1)) : -1, &a1, &a2,
                                      ix.emulate_tail | 1u /* = 1, opaque to the assembler: rs_swar.h */);
  const uint64_t before = rs_blk_count(k, c);
  *x1 = before + a1;
  *x2 = before + a2;
}

/* ---- dense rank blocks (rs_layout.h) ---- */
struct rs_dblk {
  uint32_t cnt[4], p0[4], p1[4], p2[4];
};

/* entry `idx` of the k-mer suffix table (lower, upper) */
RS_HD void rs_ftab_load(const rs_index_dev& ix, uint64_t idx, uint64_t* lo, uint64_t* hi) {
#if defined(__CUDA_ARCH__) && !defined(RS_FTAB_NO_EVICT_FIRST)
  /* one look-up per query at a random place of a table several times the size of L2: its lines are not worth keeping
   * (the eviction priority exists for 256-bit loads only: the entry's 32-byte sector is read, the entry picked) */
  unsigned long long a, b, c, d;
  asm("ld.global.nc.L2::evict_first.v4.u64 {%0,%1,%2,%3}, [%4];"
      : "=l"(a), "=l"(b), "=l"(c), "=l"(d)
      : "l"(__cvta_generic_to_global(reinterpret_cast<const uint8_t*>(ix.ftab) + (idx >> 1) * 32)));
  *lo = (idx & 1) ? c : a;
  *hi = (idx & 1) ? d : b;
#elif defined(__CUDA_ARCH__)
  const ulonglong2 e = __ldg(reinterpret_cast<const ulonglong2*>(ix.ftab) + idx);
  *lo = e.x;
  *hi = e.y;
#else
  *lo = ix.ftab[2 * idx];
  *hi = ix.ftab[2 * idx + 1];
#endif
}

RS_HD rs_dblk rs_load_dense(const rs_index_dev& ix, uint64_t b) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(ix.dense) + b * 64;
  rs_dblk k;
#if defined(__CUDA_ARCH__)
  /* two 256-bit loads (LDG.E.256 on sm_100a) instead of four 128-bit ones: every load instruction of a warp whose
   * lanes sit on different blocks costs one L1 tag look-up per lane, and K1 over dense blocks is bound by exactly
   * that (ncu: l1tex throughput 91 % with four loads per block) */
#ifdef RS_DENSE_EVICT_LAST
#define RS_DENSE_LD "ld.global.nc.L2::evict_last.v8.u32"
#else
#define RS_DENSE_LD "ld.global.nc.v8.u32"
#endif
  asm(RS_DENSE_LD " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(k.cnt[0]), "=r"(k.cnt[1]), "=r"(k.cnt[2]), "=r"(k.cnt[3]), "=r"(k.p0[0]), "=r"(k.p0[1]), "=r"(k.p0[2]), "=r"(k.p0[3])
      : "l"(__cvta_generic_to_global(p)));
  asm(RS_DENSE_LD " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(k.p1[0]), "=r"(k.p1[1]), "=r"(k.p1[2]), "=r"(k.p1[3]), "=r"(k.p2[0]), "=r"(k.p2[1]), "=r"(k.p2[2]), "=r"(k.p2[3])
      : "l"(__cvta_generic_to_global(p + 32)));
  return k;
#endif
  const rs_u4 v0 = rs_ld128(p), v1 = rs_ld128(p + 16), v2 = rs_ld128(p + 32), v3 = rs_ld128(p + 48);
  k.cnt[0] = v0.x; k.cnt[1] = v0.y; k.cnt[2] = v0.z; k.cnt[3] = v0.w;
  k.p0[0] = v1.x; k.p0[1] = v1.y; k.p0[2] = v1.z; k.p0[3] = v1.w;
  k.p1[0] = v2.x; k.p1[1] = v2.y; k.p1[2] = v2.z; k.p1[3] = v2.w;
  k.p2[0] = v3.x; k.p2[1] = v3.y; k.p2[2] = v3.z; k.p2[3] = v3.w;
  return k;
}

/* bits of plane word w that belong to the first t symbols of the block (0 <= t <= 128) */
RS_HD uint32_t rs_dense_mask(uint32_t t, uint32_t w) {
#if defined(__CUDA_ARCH__)
  /* all ones shifted right by max(32 (w + 1) - t, 0); the hardware shift gives 0 for amounts >= 32 */
  const int32_t s = (int32_t)(32u * (w + 1u)) - (int32_t)t;
  uint32_t m;
  asm("shr.b32 %0, %1, %2;" : "=r"(m) : "r"(0xFFFFFFFFu), "r"((uint32_t)(s < 0 ? 0 : s)));
  return m;
#else
  const int32_t k = (int32_t)t - (int32_t)(32u * w);
  return k <= 0 ? 0u : (k >= 32 ? 0xFFFFFFFFu : ((1u << k) - 1u));
#endif
}

RS_HD uint32_t rs_popc(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__popc(x);
#else
  return (uint32_t)__builtin_popcount(x);
#endif
}

/* occurrences of c (1..4) among the first t1 / t2 symbols of the block; c = 0 counts the '$' */
RS_HD void rs_dense_count2(const rs_dblk& k, uint32_t c, uint32_t t1, uint32_t t2, uint32_t* r1, uint32_t* r2) {
  const uint32_t cc = c - 1u;
  const uint32_t m0 = (cc & 1u) ? 0u : 0xFFFFFFFFu, m1 = (cc & 2u) ? 0u : 0xFFFFFFFFu;
  uint32_t a1 = 0, a2 = 0;
#pragma unroll
  for (uint32_t w = 0; w < 4; ++w) {
    const uint32_t hit = c ? ((k.p0[w] ^ m0) & (k.p1[w] ^ m1) & ~k.p2[w]) : k.p2[w];
    a1 += rs_popc(hit & rs_dense_mask(t1, w));
    a2 += rs_popc(hit & rs_dense_mask(t2, w));
  }
  *r1 = a1;
  *r2 = a2;
}

/* occurrences of c (0..4) before dense block b, from its header ('$' = symbols - the four bases) */
RS_HD uint64_t rs_dense_before(const rs_dblk& k, uint64_t b, uint32_t c) {
  if (c) return c == 1 ? k.cnt[0] : c == 2 ? k.cnt[1] : c == 3 ? k.cnt[2] : k.cnt[3];
  return (b << RS_DENSE_SHIFT) - ((uint64_t)k.cnt[0] + k.cnt[1] + k.cnt[2] + k.cnt[3]);
}

RS_HD uint64_t rs_dense_rank(const rs_index_dev& ix, uint32_t c, uint64_t pos) {
  const uint64_t b = pos >> RS_DENSE_SHIFT;
  const rs_dblk k = rs_load_dense(ix, b);
  uint32_t a, unused;
  rs_dense_count2(k, c, ((uint32_t)pos & (RS_DENSE_SYMS - 1u)) + 1u, 0u, &a, &unused);
  return rs_dense_before(k, b, c) + a;
}

/* symbol at pos and (when rank != null) its rank at pos, from one dense block */
RS_HD uint32_t rs_dense_symbol_rank(const rs_index_dev& ix, uint64_t pos, uint64_t* rank) {
  const uint64_t b = pos >> RS_DENSE_SHIFT;
  const rs_dblk k = rs_load_dense(ix, b);
  const uint32_t i = (uint32_t)pos & (RS_DENSE_SYMS - 1u), w = i >> 5, bit = i & 31u;
  const uint32_t q0 = w == 0 ? k.p0[0] : w == 1 ? k.p0[1] : w == 2 ? k.p0[2] : k.p0[3];
  const uint32_t q1 = w == 0 ? k.p1[0] : w == 1 ? k.p1[1] : w == 2 ? k.p1[2] : k.p1[3];
  const uint32_t q2 = w == 0 ? k.p2[0] : w == 1 ? k.p2[1] : w == 2 ? k.p2[2] : k.p2[3];
  const uint32_t sym = ((q2 >> bit) & 1u) ? 0u : 1u + ((q0 >> bit) & 1u) + 2u * ((q1 >> bit) & 1u);
  if (rank) {
    uint32_t a, unused;
    rs_dense_count2(k, sym, i + 1u, 0u, &a, &unused);
    *rank = rs_dense_before(k, b, sym) + a;
  }
  return sym;
}

/* True rank: occurrences of c in BWT[0..pos], pos < n. */
RS_HD uint64_t rs_rank(const rs_index_dev& ix, uint32_t c, uint64_t pos) {
  if (ix.dense) return rs_dense_rank(ix, c, pos);
  const uint64_t b = rs_find_block(ix, pos);
  const rs_blk k = rs_load_block(ix, b);
  uint32_t a, unused;
  rs_words_rank<RS_BLOCK_WORDS, false>(k.w, c, (int32_t)((uint32_t)(pos - k.start) + 1u), 0, &a, &unused);
  return rs_blk_count(k, c) + a;
}

/* getOcc(c, i) with the reference's semantics (i = 2^64-1 -> 0; tail clamp), rlebwt.cpp:268-301. */
RS_HD uint64_t rs_occ(const rs_index_dev& ix, uint32_t c, uint64_t i) {
  if (i == ~0ull) return 0;
  uint64_t pos = rs_ref_pos(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  return rs_rank(ix, c, pos);
}

/* updateInterval's two ranks (src/bwt/query.cpp:11-15): getOcc(c, i1) and getOcc(c, i2). */
RS_HD void rs_occ2(const rs_index_dev& ix, uint32_t c, uint64_t i1, uint64_t i2, uint64_t* r1, uint64_t* r2) {
  const bool none1 = (i1 == ~0ull), none2 = (i2 == ~0ull);
  uint64_t p1 = none1 ? 0 : rs_ref_pos(ix, i1);
  uint64_t p2 = none2 ? 0 : rs_ref_pos(ix, i2);
  if (p1 >= ix.n) p1 = ix.n - 1;
  if (p2 >= ix.n) p2 = ix.n - 1;
  if (ix.dense) {
    *r1 = none1 ? 0 : rs_dense_rank(ix, c, p1);
    *r2 = none2 ? 0 : rs_dense_rank(ix, c, p2);
    return;
  }
  const uint64_t b1 = rs_find_block(ix, p1), b2 = rs_find_block(ix, p2);
  const bool two = (b1 == b2) && p2 >= p1;
  uint64_t x1, x2;
  rs_scan2(ix, c, b1, p1, p2, two, &x1, &x2);
  if (!two) {
    uint64_t unused;
    rs_scan2(ix, c, b2, p2, p2, false, &x2, &unused);
  }
  *r1 = none1 ? 0 : x1;
  *r2 = none2 ? 0 : x2;
}

/* alphabet.h:11-32: A,C,G,T -> 1..4, every other byte -> 0 ('$'); branch-free:
 * (ch >> 1) & 3 separates A,C,T,G (0,1,2,3); the byte must also equal the letter it selects. */
RS_HD uint32_t rs_code(uint32_t ch) {
  const uint32_t sel = ((ch >> 1) & 3u) * 8u;
  const uint32_t letter = (0x47544341u >> sel) & 0xFFu; /* 'A','C','T','G' */
  const uint32_t code = (0x03040201u >> sel) & 0xFFu;   /*  1 , 2 , 4 , 3  */
  return ch == letter ? code : 0u;
}

/* A query byte that is none of "$ACGT".  The reference gives it the '$' counters of its tree (getRank,
 * alphabet.h:11-32) but compares CHARACTERS in the final run scan (rlebwt.cpp:286,293), where it matches no run:
 * getOcc(stray, i) = number of '$' before the 64-run bottom bucket that holds position i (after the tail clamp).
 * Callers keep such bytes away from findInterval except the /gt extension loop, which grows a
 * tile over whatever the window holds (service.cpp:527-530). */
RS_HD bool rs_is_stray(uint32_t ch) { return rs_code(ch) == 0 && ch != '$'; }

/* getOcc(b, i) of the reference for a stray byte b.  The bucket boundary is recovered from the blocks themselves:
 * position -> block and run inside it -> run index R; the bucket starts at run R & ~63, which lies in block
 * (R & ~63) / RS_PAYLOAD; '$' before that block from its header, plus the '$' runs in front of the boundary.
 * Plain byte loads: this path is rare and must not cost the kernels that inline it any registers. */
RS_HD uint64_t rs_block_count_before(const rs_index_dev& ix, uint64_t b, uint32_t c);
RS_HD uint64_t rs_occ_stray(const rs_index_dev& ix, uint64_t i) {
  if (i == ~0ull) return 0;
  uint64_t pos = rs_ref_pos(ix, i);
  if (pos >= ix.n) pos = ix.n - 1;
  const uint64_t b = rs_find_block(ix, pos);
  const uint8_t* runs = rs_block_ptr(ix, b) + RS_HDR_BYTES;
  uint64_t off = pos - rs_ld64(reinterpret_cast<const uint64_t*>(rs_block_ptr(ix, b)));
  uint32_t r = 0;
  while (r + 1 < RS_PAYLOAD && off >= (runs[r] & 31u)) off -= runs[r++] & 31u;
  const uint64_t first = (b * RS_PAYLOAD + r) & ~63ull; /* first run of the bucket */
  const uint64_t b2 = first / RS_PAYLOAD;
  const uint32_t o = (uint32_t)(first % RS_PAYLOAD);
  uint64_t d = rs_block_count_before(ix, b2, 0);
  runs = rs_block_ptr(ix, b2) + RS_HDR_BYTES;
  for (uint32_t t = 0; t < o; ++t)
    if ((runs[t] >> 5) == 0) d += runs[t] & 31u;
  return d;
}

/* getOcc(b, i) for any byte b (rlebwt.cpp:268-301) */
RS_HD uint64_t rs_occ_byte(const rs_index_dev& ix, uint32_t ch, uint64_t i) {
  return rs_is_stray(ch) ? rs_occ_stray(ix, i) : rs_occ(ix, rs_code(ch), i);
}

/* findInterval (src/bwt/query.cpp:24-41) for one string of len >= 1 bytes, any byte value, as the reference
 * answers it; returns the executed steps.  The batched kernel K1 is the fast path: this is for the rare queries
 * it hands over (a byte outside "$ACGT") and for callers that search a handful of strings of varying length per
 * thread (the /gt extension loop). */
RS_HD uint32_t rs_find_interval_one(const rs_index_dev& ix, const uint8_t* w, uint64_t len, uint64_t* lower, uint64_t* upper) {
  uint32_t c = rs_code(w[len - 1]);
  uint64_t lo = ix.pc[c];
  uint64_t hi = lo + (rs_is_stray(w[len - 1]) ? rs_occ_stray(ix, ix.n - 1) : ix.init_occ[c]) - 1;
  uint32_t steps = 0;
  for (uint64_t j = len - 1; j-- > 0;) {
    c = rs_code(w[j]);
    uint64_t a1, a2;
    if (rs_is_stray(w[j])) {
      a1 = rs_occ_stray(ix, lo - 1);
      a2 = rs_occ_stray(ix, hi);
    } else {
      rs_occ2(ix, c, lo - 1, hi, &a1, &a2);
    }
    ++steps;
    lo = ix.pc[c] + a1;
    hi = ix.pc[c] + a2 - 1;
    if (lo > hi) break;
  }
  *lower = lo;
  *upper = hi;
  return steps;
}

/* Symbol at pos (< n) and the rank of that symbol at pos, from one block. */
RS_HD uint32_t rs_symbol_rank(const rs_index_dev& ix, uint64_t pos, uint64_t* rank) {
  if (ix.dense) return rs_dense_symbol_rank(ix, pos, rank);
  const uint64_t b = rs_find_block(ix, pos);
  const rs_blk k = rs_load_block(ix, b);
  const int32_t t = (int32_t)((uint32_t)(pos - k.start) + 1u);
  if (!rank) {
    const uint32_t s = rs_words_symbol<RS_BLOCK_WORDS>(k.w, t);
    return s > 4u ? 0u : s;
  }
  uint32_t a;
  uint32_t sym = rs_words_symbol_rank<RS_BLOCK_WORDS>(k.w, t, &a);
  if (sym > 4u) sym = 0u; /* unreachable for a well-formed index; keeps pc[] in bounds */
  *rank = rs_blk_count(k, sym) + a;
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

/* occurrences of c before block b, from its header */
RS_HD uint64_t rs_block_count_before(const rs_index_dev& ix, uint64_t b, uint32_t c) {
  const uint8_t* p = rs_block_ptr(ix, b);
  if (c != 0) {
    const uint32_t lo = rs_ld32(reinterpret_cast<const uint32_t*>(p + 8) + (c - 1));
    const uint32_t ch = rs_ld32(reinterpret_cast<const uint32_t*>(p + 24));
    return ((uint64_t)((ch >> (8 * (c - 1))) & 0xFFu) << 32) | lo;
  }
  const rs_u4 v0 = rs_ld128(p), v1 = rs_ld128(p + 16);
  const uint64_t lo = (uint64_t)v0.z + v0.w + v1.x + v1.y;
  const uint64_t hi = (v1.z & 0xFFu) + ((v1.z >> 8) & 0xFFu) + ((v1.z >> 16) & 0xFFu) + (v1.z >> 24);
  return (((uint64_t)v0.y << 32) | v0.x) - (lo + (hi << 32));
}

/* getOccAt(c, r) (rlebwt.cpp:233-266): position of the r-th occurrence of c, 1 <= r <= tot[c].
 * The reference's select never consults m_vSum, so there is no tail behaviour to reproduce
 * (SURVEY.md A8).  r outside [1, tot[c]] yields 2^64-1. */
RS_HD uint64_t rs_occ_at(const rs_index_dev& ix, uint32_t c, uint64_t r) {
  if (r == 0 || r > ix.tot[c]) return ~0ull;
  uint64_t b = rs_ld32(ix.sel[c] + ((r - 1) >> ix.sel_shift[c]));
  /* skip blocks that end before the r-th occurrence: their successor's header says so */
  while (b + 1 < ix.nblocks && rs_block_count_before(ix, b + 1, c) < r) ++b;
  const rs_blk k = rs_load_block(ix, b);
  const uint32_t want = (uint32_t)(r - rs_blk_count(k, c));
  const uint32_t q = rs_words_select<RS_BLOCK_WORDS>(k.w, c, (int32_t)want);
  return k.start + q - 1;
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
