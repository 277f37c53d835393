/* The per-tile seed logic of find_gt_reads (src/service/service.cpp:507-711), host + device like rs_ops.h:
 * the kernel k_gt_seeds (rs_gt.cu) runs it one thread per tile, tests/emul replays it on host memory. */
#ifndef RS_GT_OPS_H
#define RS_GT_OPS_H

#include "rs_ops.h"

#ifndef RS_GT_KEEP_ALL /* also in include/readserver_b200.h */
#define RS_GT_KEEP_ALL 0u
#define RS_GT_NEED_POSTFIX 1u
#define RS_GT_NEED_PREFIX 2u
#define RS_GT_FLAG_THROW 1u
#define RS_GT_FLAG_HANG 2u
#define RS_GT_FLAG_WALK 4u
#endif

struct gt_iv {
  uint64_t lo, hi;
};

RS_HD bool gt_wide(const gt_iv& v, uint64_t max_iv) { return v.hi - v.lo + 1 > max_iv; }

/* substr(start-1, len) of a string of L bytes followed by findInterval; start is 1-based as in the reference */
RS_HD bool gt_search(const rs_index_dev& ix, const uint8_t* w, uint64_t L, uint64_t start, uint64_t len, gt_iv* v,
                          uint32_t* flag) {
  if (start < 1) {
    *flag |= RS_GT_FLAG_THROW;
    return false;
  }
  if (start - 1 + len > L) {
    *flag |= RS_GT_FLAG_HANG;
    return false;
  }
  rs_find_interval_one(ix, w + (start - 1), len, &v->lo, &v->hi);
  return true;
}

/* Tile `index` of window w (L bytes), site p (1-based): the final interval(s) whose rows the reference walks,
 * the left-hand side of its reach test for each (thr), which test applies (kind) and RS_GT_FLAG_* bits. */
RS_HD void rs_gt_tile_seeds(const rs_index_dev& ix, const uint8_t* w, uint64_t L, uint64_t p, uint64_t index,
                            uint32_t kmers, uint32_t step, uint64_t max_iv, gt_iv seed[2], uint64_t thr[2],
                            uint32_t* kind_out, uint32_t* flag_out) {
  const uint64_t start0 = (uint64_t)step * index + 1, end0 = start0 + kmers - 1; /* 1-based, inclusive */
  seed[0].lo = seed[1].lo = 1;
  seed[0].hi = seed[1].hi = 0;
  thr[0] = thr[1] = 0;
  uint32_t kind = RS_GT_KEEP_ALL, flag = 0;
  bool plain = true;
  for (uint32_t i = 0; i < kmers; ++i) plain &= rs_code(w[start0 - 1 + i]) != 0;

  if (plain) {
    gt_iv v;
    rs_find_interval_one(ix, w + (start0 - 1), kmers, &v.lo, &v.hi);
    const bool wide = gt_wide(v, max_iv);
    uint64_t start = start0, end = end0, count;
    if (p > end0) { /* tile left of the site (:522-589) */
      kind = RS_GT_NEED_POSTFIX;
      if (wide) {
        bool ok = true;
        count = 0;
        while (ok && gt_wide(v, max_iv)) { /* to the right */
          ++count;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
          ++end;
        }
        if (ok) seed[0] = v, thr[0] = p - end;
        if (ok && index != 0) { /* then one base to the left and on (not reached when the first loop hangs) */
          start = start0 - 1;
          end = end0;
          count = 1;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
          while (ok && gt_wide(v, max_iv)) {
            ++count;
            if (start > 1) --start;
            else ++end;
            ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
          }
          if (ok) seed[1] = v, thr[1] = p - end;
        }
      } else {
        seed[0] = v, thr[0] = p - end;
      }
    } else if (p < start0) { /* tile right of the site (:591-658) */
      kind = RS_GT_NEED_PREFIX;
      if (wide) {
        bool ok = true;
        count = 0;
        while (ok && gt_wide(v, max_iv)) { /* to the left */
          ++count;
          --start;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
        }
        if (ok) seed[0] = v, thr[0] = start - p;
        if (ok && end0 < L) { /* then one base to the right and on */
          start = start0;
          end = end0 + 1;
          count = 1;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
          while (ok && gt_wide(v, max_iv)) {
            ++count;
            if (end < L) ++end;
            else --start;
            ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
          }
          if (ok) seed[1] = v, thr[1] = start - p;
        }
      } else {
        seed[0] = v, thr[0] = start - p;
      }
    } else { /* tile over the site (:660-707) */
      if (wide) {
        bool ok = true;
        count = 0;
        while (ok && gt_wide(v, max_iv)) {
          ++count;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
        }
        if (ok) seed[0] = v;
        start = start0 - 1;
        count = 1;
        ok = ok && gt_search(ix, w, L, start, kmers + count, &v, &flag);
        while (ok && gt_wide(v, max_iv)) {
          ++count;
          --start;
          ok = gt_search(ix, w, L, start, kmers + count, &v, &flag);
        }
        if (ok) seed[1] = v;
      } else {
        seed[0] = v;
      }
    }
  }
  if (flag) { /* the reference does not return from this tile: none of its rows count */
    seed[0].lo = seed[1].lo = 1;
    seed[0].hi = seed[1].hi = 0;
  }
  *kind_out = kind;
  *flag_out = flag;
}

#endif
