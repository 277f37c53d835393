/* In-block rank / access / select over SGA run bytes, 32-bit SWAR.
 *
 * A thread holds the NW consecutive 32-bit payload words (4 run bytes each) of one 64-byte block
 * (NW = 9, rs_layout.h).  Everything here is branch-free so the same source compiles for the sm_100a
 * kernels and for the host, where tests/ replays it against the oracle.
 *
 * Byte-SIMD without the video intrinsics (on sm_100a __vminu4 / __vsubus4 / __vcmpeq4 expand to 5-8
 * ALU instructions each): every quantity a byte holds here is < 128 (a word spans at most 4 x 31 =
 * 124 symbols), so bit 7 of each byte is free to act as a guard bit for compares, and byte sums are
 * taken with dp4a.
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
#define RS_HIGH 0x80808080u

/* sum over bytes of a_j * b_j (unsigned) + acc */
RS_HD uint32_t rs_dot(uint32_t a, uint32_t b, uint32_t acc) {
#if defined(__CUDA_ARCH__)
  return __dp4a(a, b, acc);
#else
  for (int i = 0; i < 4; ++i) acc += ((a >> (8 * i)) & 255u) * ((b >> (8 * i)) & 255u);
  return acc;
#endif
}

RS_HD int rs_ffs(uint32_t a) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)a);
#else
  return a ? __builtin_ffs((int)a) : 0;
#endif
}

/* 8 in every byte whose run symbol equals c (cpat = c * 0x20202020), else 0 */
RS_HD uint32_t rs_w_match8(uint32_t w, uint32_t cpat) {
  const uint32_t x = (w ^ cpat) & RS_SYM_MASK;            /* 0 in the symbol bits where equal      */
#if defined(__CUDA_ARCH__)
  /* (x >> 5) + 7 per byte as ONE multiply-high-add on the FMA pipe (the ALU pipe is the busy one) */
  const uint32_t y = __umulhi(x, 0x08000000u) + 0x07070707u;
#else
  const uint32_t y = (x >> 5) + 0x07070707u;
#endif
  return ~y & 0x08080808u;                                /* bit 3 stays clear only for x == 0     */
}

/* byte j = number of symbols of run j that lie among the first tt symbols of the word;
 * p = run lengths (w & RS_LEN_MASK), 0 <= tt <= 127 */
RS_HD uint32_t rs_w_take(uint32_t p, uint32_t tt) {
  const uint32_t incl = p * RS_ONES;                      /* byte j = len_0 + .. + len_j (<= 124)  */
  const uint32_t ttr = tt * RS_ONES;
  const uint32_t d = (incl | RS_HIGH) - ttr;              /* bit 7 of byte j set iff incl_j >= tt  */
  const uint32_t ge = ((d >> 7) & RS_ONES) * 0xFFu;       /* 0xFF where incl_j >= tt               */
  const uint32_t m = (ttr & ge) | (incl & ~ge);           /* min(incl_j, tt)                       */
  return m * 0xFFFFFF01u;                                 /* m_j - m_{j-1}: no borrows, m ascends  */
}

/* index (0..3) of the first byte whose inclusive length prefix reaches tt, 1 <= tt <= word total */
RS_HD uint32_t rs_w_byte_of(uint32_t p, uint32_t tt) {
  const uint32_t incl = p * RS_ONES;
  const uint32_t d = ((incl | RS_HIGH) - tt * RS_ONES) & RS_HIGH;
  return (uint32_t)(rs_ffs(d) - 1) >> 3;
}

/* Symbols held by the NW words. */
template <int NW>
RS_HD uint32_t rs_words_total(const uint32_t (&w)[NW]) {
  uint32_t tot = 0;
#pragma unroll
  for (int j = 0; j < NW; ++j) tot = rs_dot(w[j] & RS_LEN_MASK, RS_ONES, tot);
  return tot;
}

/* Occurrences of symbol c among the first t1 (resp. t2) symbols of the words.  t may be <= 0
 * (nothing) or >= the total (everything).  One pass: whole words before the target are added with
 * one dp4a; the single word the target falls into is kept aside and resolved once after the loop. */
template <int NW, bool TWO>
RS_HD void rs_words_rank(const uint32_t (&w)[NW], uint32_t c, int32_t t1, int32_t t2, uint32_t* r1, uint32_t* r2,
                         uint32_t one = 1u) {
  const uint32_t cpat = c * 0x20202020u;
  int32_t run = 0;
  uint32_t acc1 = 0, acc2 = 0;   /* 8 x count over fully included words           */
  uint32_t bw1 = w[0], bw2 = w[0]; /* first word that is not fully included        */
  uint32_t rb1 = 0, rb2 = 0;     /* symbols before that word                       */
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const uint32_t p = w[j] & RS_LEN_MASK;
    const uint32_t cw = rs_dot(p, rs_w_match8(w[j], cpat), 0);
    const uint32_t len = rs_dot(p, RS_ONES, 0);
    const int32_t end = run + (int32_t)len;
    const uint32_t nxt = (j + 1 < NW) ? w[j + 1] : 0u;
    const uint32_t diff = nxt - w[j]; /* bw telescopes: w[0] + sum of the included differences = w[first excluded] */
#if defined(__CUDA_ARCH__)
    /* The ALU pipe (SEL / ISETP / LOP3 / IADD3, 16 lanes per cycle) is K1's busiest unit (ncu: 70 % against
     * 10 % for the FMA pipe), so the three conditional updates per target are predicated multiply-adds, which
     * issue on the FMA pipe; `one` is a 1 the assembler cannot see through (it would turn them back into adds). */
    asm("{\n\t.reg .pred q;\n\tsetp.le.s32 q, %3, %4;\n\t@q mad.lo.u32 %0, %5, %8, %0;\n\t@q mad.lo.u32 %1, %6, %8, %1;\n\t"
        "@q mad.lo.u32 %2, %7, %8, %2;\n\t}"
        : "+r"(acc1), "+r"(rb1), "+r"(bw1)
        : "r"(end), "r"(t1), "r"(cw), "r"(len), "r"(diff), "r"(one));
    if (TWO)
      asm("{\n\t.reg .pred q;\n\tsetp.le.s32 q, %3, %4;\n\t@q mad.lo.u32 %0, %5, %8, %0;\n\t@q mad.lo.u32 %1, %6, %8, %1;\n\t"
          "@q mad.lo.u32 %2, %7, %8, %2;\n\t}"
          : "+r"(acc2), "+r"(rb2), "+r"(bw2)
          : "r"(end), "r"(t2), "r"(cw), "r"(len), "r"(diff), "r"(one));
#else
    (void)one;
    if (end <= t1) { acc1 += cw; rb1 += len; bw1 += diff; }
    if (TWO && end <= t2) { acc2 += cw; rb2 += len; bw2 += diff; }
#endif
    run = end;
  }
  {
    int32_t tt = t1 - (int32_t)rb1;
    tt = tt < 0 ? 0 : (tt > 127 ? 127 : tt);
    acc1 = rs_dot(rs_w_take(bw1 & RS_LEN_MASK, (uint32_t)tt), rs_w_match8(bw1, cpat), acc1);
    *r1 = acc1 >> 3;
  }
  if (TWO) {
    int32_t tt = t2 - (int32_t)rb2;
    tt = tt < 0 ? 0 : (tt > 127 ? 127 : tt);
    acc2 = rs_dot(rs_w_take(bw2 & RS_LEN_MASK, (uint32_t)tt), rs_w_match8(bw2, cpat), acc2);
    *r2 = acc2 >> 3;
  }
}

/* Symbol (0..4) of the t-th symbol of the words, 1 <= t <= total; 0xFF when t is outside. */
template <int NW>
RS_HD uint32_t rs_words_symbol(const uint32_t (&w)[NW], int32_t t) {
  int32_t run = 0, rb = 0;
  uint32_t bw = w[0];
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const int32_t end = run + (int32_t)rs_dot(w[j] & RS_LEN_MASK, RS_ONES, 0);
    const uint32_t nxt = (j + 1 < NW) ? w[j + 1] : 0u;
    if (end < t) { bw = nxt; rb = end; }   /* word lies wholly before the t-th symbol */
    run = end;
  }
  if (t < 1 || t > run) return 0xFFu;
  const uint32_t byte = rs_w_byte_of(bw & RS_LEN_MASK, (uint32_t)(t - rb));
  return (bw >> (8 * byte + 5)) & 7u;
}

/* Symbol of the t-th symbol of the words (1 <= t <= total; 0xFF when outside) and, in *rank, the occurrences
 * of that symbol among the first t symbols: one pass over the run lengths finds the word, one pass counts the
 * matches in the words before it (the LF step of extractPrefix needs both, rlebwt.cpp:202-227 + 268-301). */
template <int NW>
RS_HD uint32_t rs_words_symbol_rank(const uint32_t (&w)[NW], int32_t t, uint32_t* rank) {
  int32_t run = 0, rb = 0;
  uint32_t bw = w[0];
  int32_t jw = 0; /* index of the word holding the t-th symbol */
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const int32_t end = run + (int32_t)rs_dot(w[j] & RS_LEN_MASK, RS_ONES, 0);
    const uint32_t nxt = (j + 1 < NW) ? w[j + 1] : 0u;
    if (end < t) { bw = nxt; rb = end; jw = j + 1; }
    run = end;
  }
  *rank = 0;
  if (t < 1 || t > run) return 0xFFu;
  const uint32_t p = bw & RS_LEN_MASK;
  const uint32_t byte = rs_w_byte_of(p, (uint32_t)(t - rb));
  const uint32_t sym = (bw >> (8 * byte + 5)) & 7u;
  const uint32_t cpat = sym * 0x20202020u;
  uint32_t acc = 0;
#pragma unroll
  for (int j = 0; j < NW; ++j)
    if (j < jw) acc = rs_dot(w[j] & RS_LEN_MASK, rs_w_match8(w[j], cpat), acc);
  acc = rs_dot(rs_w_take(p, (uint32_t)(t - rb)), rs_w_match8(bw, cpat), acc);
  *rank = acc >> 3;
  return sym;
}

/* Occurrences of c in all the words. */
template <int NW>
RS_HD uint32_t rs_words_count_all(const uint32_t (&w)[NW], uint32_t c) {
  const uint32_t cpat = c * 0x20202020u;
  uint32_t acc = 0;
#pragma unroll
  for (int j = 0; j < NW; ++j) acc = rs_dot(w[j] & RS_LEN_MASK, rs_w_match8(w[j], cpat), acc);
  return acc >> 3;
}

/* Select inside the words: 1-based position (counted in symbols from the first word) of the r-th
 * occurrence of c, 1 <= r <= rs_words_count_all; 0 when r is outside. */
template <int NW>
RS_HD uint32_t rs_words_select(const uint32_t (&w)[NW], uint32_t c, int32_t r) {
  const uint32_t cpat = c * 0x20202020u;
  int32_t run = 0, rb = 0, seen = 0, sb = 0;
  uint32_t bw = w[0];
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const uint32_t p = w[j] & RS_LEN_MASK;
    const int32_t cend = seen + (int32_t)(rs_dot(p, rs_w_match8(w[j], cpat), 0) >> 3);
    const int32_t end = run + (int32_t)rs_dot(p, RS_ONES, 0);
    const uint32_t nxt = (j + 1 < NW) ? w[j + 1] : 0u;
    if (cend < r) { bw = nxt; rb = end; sb = cend; }   /* word holds none of the r-th occurrence */
    run = end;
    seen = cend;
  }
  if (r < 1 || r > seen) return 0u;
  /* inside bw: lengths of the matching runs only, then the byte whose matching prefix reaches rk */
  const uint32_t p = bw & RS_LEN_MASK;
  const uint32_t m8 = rs_w_match8(bw, cpat);
  const uint32_t mp = p & ((m8 >> 3) * 0xFFu);
  const uint32_t rk = (uint32_t)(r - sb);
  const uint32_t byte = rs_w_byte_of(mp, rk);
  const uint32_t before = ((p * RS_ONES - p) >> (8 * byte)) & 255u;      /* symbols before that run in the word  */
  const uint32_t mbefore = ((mp * RS_ONES - mp) >> (8 * byte)) & 255u;   /* matching symbols before that run     */
  return (uint32_t)rb + before + (rk - mbefore);
}

#endif
