/* Per-lane arithmetic of the in-block rank / access / select over SGA run bytes.
 *
 * One lane owns 16 bytes (four 32-bit words) of a 128-byte block.  Everything here is pure 32-bit
 * integer SWAR over four run bytes at a time, so the same source compiles for the sm_100a kernels
 * (video-SIMD + dp4a instructions) and for the host (portable fall-backs) where tests/ replays it
 * lane by lane against the oracle.  The cross-lane glue (scan / reduce / broadcast) lives with the
 * caller: warp shuffles on the device.
 *
 * Replaces the byte-at-a-time run scans of the reference: RLEBWT::getOcc (src/bwt/rlebwt.cpp:281-298),
 * getChar (:213-225) and getOccAt (:247-264).
 */
#ifndef RS_SWAR_H
#define RS_SWAR_H

#include <stdint.h>

#if defined(__CUDACC__)
#define RS_HD __host__ __device__ __forceinline__
#else
#define RS_HD static inline
#endif

#define RS_LEN_MASK 0x1F1F1F1Fu
#define RS_SYM_MASK 0xE0E0E0E0u
#define RS_ONES 0x01010101u

/* ---- byte-SIMD primitives ---- */
#if defined(__CUDA_ARCH__)
RS_HD uint32_t rs_vsubus4(uint32_t a, uint32_t b) { return __vsubus4(a, b); }
RS_HD uint32_t rs_vminu4(uint32_t a, uint32_t b) { return __vminu4(a, b); }
RS_HD uint32_t rs_vcmpeq4(uint32_t a, uint32_t b) { return __vcmpeq4(a, b); }
RS_HD uint32_t rs_vcmpgeu4(uint32_t a, uint32_t b) { return __vcmpgeu4(a, b); }
RS_HD uint32_t rs_bytesum(uint32_t a, uint32_t acc) { return __dp4a(a, RS_ONES, acc); }
RS_HD int rs_ffs(uint32_t a) { return __ffs((int)a); }
#else
RS_HD uint32_t rs_vsubus4(uint32_t a, uint32_t b) {
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) {
    uint32_t x = (a >> (8 * i)) & 255u, y = (b >> (8 * i)) & 255u;
    r |= (x > y ? x - y : 0u) << (8 * i);
  }
  return r;
}
RS_HD uint32_t rs_vminu4(uint32_t a, uint32_t b) {
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) {
    uint32_t x = (a >> (8 * i)) & 255u, y = (b >> (8 * i)) & 255u;
    r |= (x < y ? x : y) << (8 * i);
  }
  return r;
}
RS_HD uint32_t rs_vcmpeq4(uint32_t a, uint32_t b) {
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i)
    if (((a >> (8 * i)) & 255u) == ((b >> (8 * i)) & 255u)) r |= 255u << (8 * i);
  return r;
}
RS_HD uint32_t rs_vcmpgeu4(uint32_t a, uint32_t b) {
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i)
    if (((a >> (8 * i)) & 255u) >= ((b >> (8 * i)) & 255u)) r |= 255u << (8 * i);
  return r;
}
RS_HD uint32_t rs_bytesum(uint32_t a, uint32_t acc) {
  return acc + (a & 255u) + ((a >> 8) & 255u) + ((a >> 16) & 255u) + (a >> 24);
}
RS_HD int rs_ffs(uint32_t a) { return a ? __builtin_ffs((int)a) : 0; }
#endif

/* Decoded view of one lane's 16 bytes. */
typedef struct rs_lane {
  uint32_t w[4];    /* run bytes (header bytes zeroed)                     */
  uint32_t incl[4]; /* byte j = sum of lengths of bytes 0..j of the word   */
  uint32_t base[4]; /* symbols in the lane before word k                    */
  uint32_t total;   /* symbols in the lane                                   */
} rs_lane;

/* g = lane position in its group (0..7); x,y,z,w = the lane's 16 bytes of the block.
 * Lanes 0,1 hold header only, lane 2 holds cnt[T] in x,y and run bytes in z,w. */
RS_HD void rs_lane_prepare(uint32_t x, uint32_t y, uint32_t z, uint32_t w, int g, rs_lane* d) {
  d->w[0] = g >= 3 ? x : 0u;
  d->w[1] = g >= 3 ? y : 0u;
  d->w[2] = g >= 2 ? z : 0u;
  d->w[3] = g >= 2 ? w : 0u;
  uint32_t run = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t p = d->w[k] & RS_LEN_MASK;
    d->incl[k] = p * RS_ONES; /* no byte overflows: 4 x 31 = 124 */
    d->base[k] = run;
    run += d->incl[k] >> 24;
  }
  d->total = run;
}

/* Occurrences of symbol c among the first t symbols of the lane (t may be <= 0 or > total). */
RS_HD uint32_t rs_lane_count(const rs_lane* d, uint32_t c, int32_t t) {
  const uint32_t cpat = c * 0x20202020u;
  uint32_t acc = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int32_t tk = t - (int32_t)d->base[k];
    tk = tk < 0 ? 0 : (tk > 255 ? 255 : tk);
    const uint32_t p = d->w[k] & RS_LEN_MASK;
    const uint32_t excl = d->incl[k] - p;
    const uint32_t take = rs_vminu4(p, rs_vsubus4((uint32_t)tk * RS_ONES, excl));
    const uint32_t match = rs_vcmpeq4(d->w[k] & RS_SYM_MASK, cpat);
    acc = rs_bytesum(take & match, acc);
  }
  return acc;
}

/* Occurrences of symbol c in the whole lane. */
RS_HD uint32_t rs_lane_count_all(const rs_lane* d, uint32_t c) {
  const uint32_t cpat = c * 0x20202020u;
  uint32_t acc = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t match = rs_vcmpeq4(d->w[k] & RS_SYM_MASK, cpat);
    acc = rs_bytesum(d->w[k] & RS_LEN_MASK & match, acc);
  }
  return acc;
}

/* Symbol (0..4) of the t-th symbol of the lane, 1 <= t <= total; 0xFF if t is outside the lane. */
RS_HD uint32_t rs_lane_symbol(const rs_lane* d, int32_t t) {
  uint32_t sym = 0xFFu;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int32_t tk = t - (int32_t)d->base[k];
    const int32_t wt = (int32_t)(d->incl[k] >> 24);
    if (tk >= 1 && tk <= wt) {
      const uint32_t ge = rs_vcmpgeu4(d->incl[k], (uint32_t)tk * RS_ONES) & 0x80808080u;
      const int byte = (rs_ffs(ge) - 1) >> 3;
      sym = (d->w[k] >> (8 * byte + 5)) & 7u;
    }
  }
  return sym;
}

/* Select inside a lane: position (1-based count of symbols from the lane start) of the r-th
 * occurrence of c in the lane, 1 <= r <= rs_lane_count_all; 0 if r is outside. */
RS_HD uint32_t rs_lane_select(const rs_lane* d, uint32_t c, int32_t r) {
  const uint32_t cpat = c * 0x20202020u;
  uint32_t pos = 0;
  int32_t seen = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t match = rs_vcmpeq4(d->w[k] & RS_SYM_MASK, cpat);
    const uint32_t p = d->w[k] & RS_LEN_MASK;
    const uint32_t mp = p & match;           /* lengths of matching runs               */
    const uint32_t mincl = mp * RS_ONES;     /* inclusive prefix of matching lengths   */
    const int32_t mt = (int32_t)(mincl >> 24);
    const int32_t rk = r - seen;
    if (rk >= 1 && rk <= mt) {
      const uint32_t ge = rs_vcmpgeu4(mincl, (uint32_t)rk * RS_ONES) & 0x80808080u;
      const int byte = (rs_ffs(ge) - 1) >> 3;
      /* symbols before that run inside the word + offset inside the run */
      const uint32_t before = ((d->incl[k] - p) >> (8 * byte)) & 255u;
      const uint32_t mbefore = ((mincl - mp) >> (8 * byte)) & 255u;
      pos = d->base[k] + before + ((uint32_t)rk - mbefore);
    }
    seen += mt;
  }
  return pos;
}

#endif
