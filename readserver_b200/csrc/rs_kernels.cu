/* sm_100a kernels of the FM-index hot path.  One thread per query / walker (rs_ops.h); the lanes of
 * a warp always run the same code on different queries.
 *
 *   k_find_interval   findInterval            src/bwt/query.cpp:24-41   (K1)
 *   k_extract_prefix  extractPrefix           src/bwt/query.cpp:43-63   (K2)
 *   k_extract_postfix extractPostfix          src/bwt/query.cpp:65-85
 *   k_occ / k_char / k_occ_at / k_getf        src/bwt/rlebwt.cpp:202-314
 */
#include "rs_kernels.h"

#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "rs_ops.h"

#define RS_FULL 0xFFFFFFFFu

__device__ __forceinline__ uint32_t rs_ascii(uint32_t code) { return (0x5447434124ull >> (8 * code)) & 0xFFu; }

__global__ void __launch_bounds__(RS_THREADS) k_occ(const __grid_constant__ rs_index_dev ix, const uint8_t* __restrict__ syms,
                                                    const uint64_t* __restrict__ idx, uint64_t n,
                                                    uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = rs_occ_byte(ix, syms[i], idx[i]);
}

__global__ void __launch_bounds__(RS_THREADS) k_char(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ idx,
                                                     uint64_t n, uint8_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = (uint8_t)rs_ascii(rs_char(ix, idx[i]));
}

__global__ void __launch_bounds__(RS_THREADS) k_occ_at(const __grid_constant__ rs_index_dev ix, const uint8_t* __restrict__ syms,
                                                       const uint64_t* __restrict__ cnt, uint64_t n,
                                                       uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = rs_occ_at(ix, rs_code(syms[i]), cnt[i]);
}

/* getF (rlebwt.cpp:307-314): last symbol whose C[] value is <= index. One thread per item. */
__global__ void k_getf(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ idx, uint64_t n,
                       uint8_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t v = idx[i];
  uint32_t ci = 0;
  while (ci < RS_NSYM && ix.pc[ci] <= v) ++ci;
  out[i] = (uint8_t)rs_ascii(ci - 1);
}

/* Pre-pass of K1 for fixed-length queries of k <= 31 characters: the searched string (the query, or its
 * reverse complement when rc) as a base-4 number, first character most significant, A,C,G,T = 0..3;
 * bit 63 = the query holds a byte outside ACGT.  A block stages its 256 queries through shared memory
 * so that the ASCII batch is read with coalesced loads.  The result is parked in lower[] (each K1
 * thread reads lower[q] before it writes it). */
#define RS_PACK_MAX_K 31
__global__ void __launch_bounds__(RS_THREADS) k_pack_kmers(const uint8_t* __restrict__ text, uint32_t k, uint64_t Q,
                                                          int rc, uint64_t* __restrict__ packed) {
  __shared__ uint8_t sm[RS_THREADS * RS_PACK_MAX_K];
  const uint64_t q0 = (uint64_t)blockIdx.x * RS_THREADS;
  const uint32_t nq = (uint32_t)(Q - q0 < RS_THREADS ? Q - q0 : RS_THREADS);
  const uint32_t nbytes = nq * k;
  const uint8_t* src = text + q0 * k;
  const uint32_t head = (uint32_t)((4 - ((uintptr_t)src & 3)) & 3); /* bytes before the first aligned word */
  for (uint32_t i = threadIdx.x; i < head && i < nbytes; i += RS_THREADS) sm[i] = src[i];
  if (nbytes > head) {
    /* aligned words are stored to shared memory byte by byte: sm has no alignment relation to src */
    const uint32_t nwords = (nbytes - head) / 4;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(src + head);
    for (uint32_t i = threadIdx.x; i < nwords; i += RS_THREADS) {
      const uint32_t v = __ldg(w + i);
      uint8_t* d = sm + head + 4 * i;
      d[0] = (uint8_t)v;
      d[1] = (uint8_t)(v >> 8);
      d[2] = (uint8_t)(v >> 16);
      d[3] = (uint8_t)(v >> 24);
    }
    for (uint32_t i = head + 4 * nwords + threadIdx.x; i < nbytes; i += RS_THREADS) sm[i] = src[i];
  }
  __syncthreads();
  if (threadIdx.x >= nq) return;
  const uint8_t* s = sm + threadIdx.x * k;
  uint64_t v = 0;
  uint32_t bad = 0;
  for (uint32_t i = 0; i < k; ++i) {
    const uint32_t code = rs_code(rc ? s[k - 1 - i] : s[i]);
    bad |= (code == 0);
    const uint32_t two = (code - 1u) & 3u;
    v = (v << 2) | (rc ? 3u - two : two);
  }
  packed[q0 + threadIdx.x] = v | ((uint64_t)(bad != 0) << 63);
}

/* K1.  Each thread owns one query at a time and all lanes of a warp perform one block scan per
 * loop trip, whatever query they are on: a thread whose query ends (all characters consumed, or the
 * interval went empty: query.cpp:35-37) picks up its next query on the following trip, so lanes never
 * idle behind a longer neighbour.  A step (updateInterval, query.cpp:11-15) whose two positions share
 * a block takes one trip; otherwise the second position is scanned on the next trip.
 *   mode 0: write lower/upper.   mode 1: write counts (count_reads, service.cpp:279-315), strand as given.
 *   mode 2: add counts into the owning rank's buffer (rs_emit below). */

/* Where a finished query's answer goes.  mode 0: lower/upper; mode 1: count into lower[];
 * mode 2 (packed queries only): the count is added into the accumulation buffer of the rank that
 * owns the k-mer -- its first four bases as a base-4 number, modulo the number of ranks -- directly
 * through NVLink peer memory: K1 and the reduce-scatter of the per-shard counts are one kernel
 * (replaces the front end's fan-in sum, src/service/server.cpp:184-197, 404-410). */
__device__ __forceinline__ uint32_t rs_owner_of(uint64_t pk, uint32_t k, bool rc, uint32_t world) {
  uint32_t prefix;
  if (!rc) {
    prefix = k >= 4 ? (uint32_t)(pk >> (2 * (k - 4))) & 0xFFu : (uint32_t)pk & 0xFFu;
  } else { /* pk holds the reverse complement: the query's first bases are its last ones, complemented */
    const uint32_t t = ~(uint32_t)pk & 0xFFu;
    if (k >= 4) {
      prefix = ((t & 3u) << 6) | (((t >> 2) & 3u) << 4) | (((t >> 4) & 3u) << 2) | (t >> 6);
    } else { /* fewer than four bases: reverse the k complemented digits only */
      prefix = 0;
      for (uint32_t a = 0; a < k; ++a) prefix = (prefix << 2) | ((t >> (2 * a)) & 3u);
    }
  }
  return prefix % world;
}

__device__ __forceinline__ void rs_emit(int mode, uint64_t* __restrict__ lower, uint64_t* __restrict__ upper,
                                        const rs_peers& peers, uint64_t q, uint64_t lo, uint64_t hi, uint64_t pk,
                                        uint32_t k, bool rc) {
  if (mode == 0) {
    lower[q] = lo;
    upper[q] = hi;
  } else if (mode == 4) { /* shard set (rs_pop.cu): first row and number of rows, 12 bytes per shard-query */
    const uint64_t cnt = hi >= lo ? hi - lo + 1 : 0;
#ifndef RS_K1_NO_STREAM_OUT /* written once, read once by another kernel: not worth an L2 line the rank blocks could hold */
    __stcs(reinterpret_cast<unsigned long long*>(lower) + q, (unsigned long long)lo);
    __stcs(reinterpret_cast<uint32_t*>(upper) + q, cnt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)cnt);
#else
    lower[q] = lo;
    reinterpret_cast<uint32_t*>(upper)[q] = cnt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)cnt;
#endif
  } else {
    const uint64_t cnt = hi >= lo ? hi - lo + 1 : 0;
    if (mode == 1) lower[q] = cnt;
    else if (cnt) atomicAdd(reinterpret_cast<unsigned long long*>(peers.buf[rs_owner_of(pk, k, rc, peers.world)]) + q, cnt);
  }
}

#ifndef RS_K1_RESTRICT
#define RS_K1_RESTRICT
#endif
#ifndef RS_K1_BLOCKS
#define RS_K1_BLOCKS 5 /* resident CTAs per SM the register allocation must allow (4: 1532, 5: 1573, 6: 1477 M q/s) */
#endif
template <bool FTAB, bool PACKED>
__global__ void __launch_bounds__(RS_THREADS, RS_K1_BLOCKS)
    k_find_interval(const __grid_constant__ rs_index_dev ix, const uint8_t* __restrict__ text,
                    const uint64_t* __restrict__ offs, uint32_t k, uint64_t Q, const uint64_t* pkeys, uint64_t* RS_K1_RESTRICT lower,
                    uint64_t* __restrict__ upper, int mode, int strand, unsigned long long* __restrict__ stats,
                    const __grid_constant__ rs_peers peers) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  uint64_t next_q = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x;
  uint64_t lo = 0, hi = 0;
  const uint8_t* w = text; /* the query's characters */
  int32_t j = -1;          /* index of the next character in search order */
  uint32_t len = 0;
  bool active = false;
  const bool rc = (mode >= 1 && strand == 1);
  uint32_t nsteps = 0; /* updateInterval calls executed by this thread (2 ranks each) */
  uint32_t nscans = 0; /* 64-byte run blocks read (one per trip) */
  bool pending = false; /* the second position of the step in hand is still to be scanned */
  uint32_t c = 0, b2 = 0;
  uint64_t r1 = 0;
  uint64_t pk = 0; /* PACKED: the searched string as k_pack_kmers left it in lower[q] */

  while (true) {
    if (PACKED && !active && next_q < Q) {
      const uint64_t q = next_q;
      next_q += stride;
      pk = pkeys[q]; /* PACKED: k_pack_kmers' output; parked in lower[] itself unless the caller keeps it apart */
      len = k;
      active = true;
      if (pk >> 63) { /* a byte outside ACGT */
        if (mode == 1) lower[q] = 0; /* service.cpp:299-301; mode 2: nothing to add */
        if (mode == 4) { lower[q] = 1; reinterpret_cast<uint32_t*>(upper)[q] = 0; } /* count_reads: 0 rows */
        /* mode 0: left flagged for the tail loop below, which searches it as the reference would */
        active = false;
      } else {
        const uint32_t c0 = (uint32_t)(pk & 3u) + 1u; /* initInterval (query.cpp:18-21) on the last character */
        lo = ix.pc[c0];
        hi = lo + ix.init_occ[c0] - 1;
        j = (int32_t)len - 2;
        if (FTAB && len >= ix.ftab_k) {
          const uint32_t t = ix.ftab_k;
          const ulonglong2 e = __ldg(reinterpret_cast<const ulonglong2*>(ix.ftab) + (pk & ((1ull << (2 * t)) - 1)));
          lo = e.x;
          hi = e.y;
          j = (int32_t)len - (int32_t)t - 1;
          if (lo > hi) j = -1; /* the search had already stopped inside the suffix (query.cpp:35-37) */
        }
        if (j < 0) {
          rs_emit(mode, lower, upper, peers, q, lo, hi, pk, k, rc);
          active = false;
        }
      }
    }
    if (!PACKED && !active && next_q < Q) {
      const uint64_t q = next_q;
      next_q += stride;
      uint64_t base;
      if (offs) {
        base = offs[q];
        len = (uint32_t)(offs[q + 1] - base);
      } else {
        base = q * (uint64_t)k;
        len = k;
      }
      w = text + base;
      active = true;
      bool bad = (len == 0);
      if (mode == 1) /* any byte outside ACGT -> count 0 (service.cpp:299-301) */
        for (uint32_t t = 0; t < len; ++t) bad |= (rs_code(w[t]) == 0);
      if (bad) {
        if (mode == 1) lower[q] = 0;
        else { lower[q] = 1; upper[q] = 0; }
        active = false;
      } else if (mode == 0 && rs_is_stray(w[len - 1])) {
        lower[q] = 1ull << 63; /* searched whole by the tail loop below */
        active = false;
      } else {
        /* initInterval (query.cpp:18-21) on the last character of the searched string */
        const uint32_t c0 = rc ? 5u - rs_code(w[0]) : rs_code(w[len - 1]);
        lo = ix.pc[c0];
        hi = lo + ix.init_occ[c0] - 1;
        j = (int32_t)len - 2;
        if (FTAB && len >= ix.ftab_k) {
          /* the first ftab_k - 1 steps are a table lookup when the last ftab_k characters are all ACGT */
          const uint32_t t = ix.ftab_k;
          uint32_t idx = 0, ok = 1;
          for (uint32_t a = 0; a < t; ++a) {
            const uint32_t cc = rc ? 5u - rs_code(w[t - 1 - a]) : rs_code(w[len - t + a]);
            ok &= (cc - 1u) < 4u;
            idx = idx * 4u + ((cc - 1u) & 3u);
          }
          if (ok) {
            const ulonglong2 e = __ldg(reinterpret_cast<const ulonglong2*>(ix.ftab) + idx);
            lo = e.x;
            hi = e.y;
            j = (int32_t)len - (int32_t)t - 1;
            if (lo > hi) j = -1; /* the search had already stopped inside the suffix (query.cpp:35-37) */
          }
        }
        if (j < 0) {
          if (mode == 1) lower[q] = hi >= lo ? hi - lo + 1 : 0;
          else { lower[q] = lo; upper[q] = hi; }
          active = false;
        }
      }
    }
    if (__all_sync(RS_FULL, !active)) {
      if (__all_sync(RS_FULL, next_q >= Q)) break;
      continue;
    }
    if (active) {
      /* positions of the step in hand (recomputed on the second trip of a split step: cheaper than
       * keeping them in registers across trips) */
      const bool none1 = (lo == 0), none2 = (hi == ~0ull); /* getOcc(b, 2^64-1) = 0 in the reference */
      uint64_t p1 = none1 ? 0 : rs_ref_pos(ix, lo - 1);
      uint64_t p2 = none2 ? 0 : rs_ref_pos(ix, hi);
      if (p1 >= ix.n) p1 = ix.n - 1;
      if (p2 >= ix.n) p2 = ix.n - 1;
      uint64_t pos, blk;
      bool two = false;
      const bool was_pending = pending;
      if (!pending) {
        if (PACKED) c = (uint32_t)((pk >> (2 * (len - 1 - (uint32_t)j))) & 3u) + 1u;
        /* reverse complement: revcomp(w)[j] is the complement of w[len-1-j] */
        else c = rc ? 5u - rs_code(w[len - 1 - j]) : rs_code(w[j]);
        if (!PACKED && mode == 0 && c == 0 && w[j] != '$') { /* a stray byte: the tail loop searches this query */
          lower[next_q - stride] = 1ull << 63;
          active = false;
          continue;
        }
        const uint64_t b1 = rs_find_block(ix, p1);
        b2 = (uint32_t)rs_find_block(ix, p2);
        two = (b1 == b2) && p2 >= p1;
        pos = p1;
        blk = b1;
        ++nsteps;
      } else {
        pos = p2;
        blk = b2;
      }
      uint64_t x1, x2;
      rs_scan2(ix, c, blk, pos, p2, two, &x1, &x2);
#ifndef RS_K1_NO_SCAN_COUNT
      ++nscans;
#endif
      uint64_t r2 = 0;
      if (was_pending) {
        r2 = none2 ? 0 : x1;
        pending = false;
      } else {
        r1 = none1 ? 0 : x1;
        r2 = none2 ? 0 : x2;
        pending = !two;
      }
      if (!pending) {
        const uint64_t pb = ix.pc[c];
        lo = pb + r1;
        hi = pb + r2 - 1;
        --j;
        if (lo > hi || j < 0) {
          rs_emit(mode, lower, upper, peers, next_q - stride, lo, hi, pk, k, rc);
          active = false;
        }
      }
    }
  }
  if (stats && nsteps) {
    atomicAdd(stats, (unsigned long long)nsteps);
    atomicAdd(stats + 2, (unsigned long long)nscans);
  }
}

/* K1 over the dense rank blocks (rs_layout.h), packed queries.  A step (updateInterval, query.cpp:11-15) is one trip:
 * the blocks of lower-1 and upper are addressed by position alone, so both loads issue at once (no directory word, no
 * second trip for a step that spans two blocks) and each rank is a few logic + popcount instructions.  Lanes whose
 * query has ended wait until RS_REFILL_MIN of the warp's lanes are idle: the set-up path (key, suffix table, emit) then
 * runs for several lanes at once instead of for one lane on almost every trip. */
#ifndef RS_K1D_BLOCKS
#define RS_K1D_BLOCKS 5 /* resident CTAs per SM (4: 137, 5: 129, 6: 165 ms per configs[2] step, spills at 6) */
#endif
#ifndef RS_REFILL_MIN
#define RS_REFILL_MIN 6
#endif
/* SMALL: the shard holds fewer than 2^31 symbols, so rows and counts live in 32-bit registers (half the ALU work of
 * the position arithmetic; 2^32-1 stands for the reference's 2^64-1 = "upper bound before row 0"). */
template <bool FTAB, bool SMALL>
__global__ void __launch_bounds__(RS_THREADS, RS_K1D_BLOCKS)
    k_find_interval_dense(const __grid_constant__ rs_index_dev ix, uint32_t k, uint64_t Q, const uint64_t* pkeys,
                          uint64_t* lower, uint64_t* __restrict__ upper, int mode, int strand,
                          unsigned long long* __restrict__ stats, const __grid_constant__ rs_peers peers,
                          unsigned long long* __restrict__ ticket) {
  typedef typename std::conditional<SMALL, uint32_t, uint64_t>::type P;
  const P NONE = (P)~(P)0;
  const uint32_t lane = threadIdx.x & 31u;
  uint64_t q = 0, pk = 0;
  P lo = 0, hi = 0;
  int32_t j = -1;
  bool active = false, exhausted = false;
  const bool rc = (mode >= 1 && mode != 4 && strand == 1);
  uint32_t nsteps = 0, nsplit = 0;
  /* the reference's tail rule (rs_ref_pos) and the clamp to the last row, in the position type */
  const P tail_from = ix.tail_from > (uint64_t)NONE ? NONE : (P)ix.tail_from;
  const P tail_pos = (P)(ix.clamp_c - 1), last = (P)(ix.n - 1);
  while (true) {
    const uint32_t idle = __ballot_sync(RS_FULL, !active);
    if (idle == RS_FULL || __popc(idle) >= RS_REFILL_MIN) {
      /* Queries are handed out from one counter, a warp taking as many as it has idle lanes: lanes of a warp (and
       * warps of the grid) stay busy until the batch runs dry, however unevenly the searches end.  With a fixed
       * share of the batch per thread a warp ran until its slowest lane was done: a third of the lane-trips idle. */
      if (!exhausted) {
        const int leader = __ffs((int)idle) - 1;
        unsigned long long base = 0;
        if ((int)lane == leader) base = atomicAdd(ticket, (unsigned long long)__popc(idle));
        base = __shfl_sync(RS_FULL, base, leader);
        if (base >= Q) exhausted = true;
        if (!active && !exhausted) {
          q = base + (uint64_t)__popc(idle & ((1u << lane) - 1u));
          if (q < Q) {
            pk = pkeys[q];
            active = true;
            if (pk >> 63) { /* a byte outside ACGT */
              if (mode == 1) lower[q] = 0; /* service.cpp:299-301; mode 2: nothing to add */
              if (mode == 4) { lower[q] = 1; reinterpret_cast<uint32_t*>(upper)[q] = 0; }
              active = false; /* mode 0: left flagged for k_search_flagged */
            } else {
              const uint32_t c0 = (uint32_t)(pk & 3u) + 1u; /* initInterval (query.cpp:18-21) on the last character */
              lo = (P)ix.pc[c0];
              hi = (P)(ix.pc[c0] + ix.init_occ[c0] - 1);
              j = (int32_t)k - 2;
              if (FTAB && k >= ix.ftab_k) {
                const uint32_t t = ix.ftab_k;
                ulonglong2 e;
                rs_ftab_load(ix, pk & ((1ull << (2 * t)) - 1), (uint64_t*)&e.x, (uint64_t*)&e.y);
                lo = (P)e.x;
                hi = (P)e.y; /* 2^64-1 becomes NONE */
                j = (int32_t)k - (int32_t)t - 1;
                if (e.x > e.y) j = -1; /* the search had already stopped inside the suffix (query.cpp:35-37) */
              }
              if (j < 0) {
                rs_emit(mode, lower, upper, peers, q, (uint64_t)lo, hi == NONE ? ~0ull : (uint64_t)hi, pk, k, rc);
                active = false;
              }
            }
          }
        }
      }
      if (__all_sync(RS_FULL, !active)) {
        if (exhausted) break;
        continue;
      }
    }
    if (active) {
      const uint32_t c = (uint32_t)((pk >> (2 * (k - 1 - (uint32_t)j))) & 3u) + 1u;
      const bool none1 = (lo == 0), none2 = (hi == NONE); /* getOcc(b, 2^64-1) = 0 in the reference */
      P p1 = lo - 1, p2 = hi;
      p1 = p1 >= tail_from ? tail_pos : p1;
      p2 = p2 >= tail_from ? tail_pos : p2;
      p1 = none1 ? 0 : (p1 > last ? last : p1);
      p2 = none2 ? 0 : (p2 > last ? last : p2);
      const P b1 = p1 >> RS_DENSE_SHIFT, b2 = p2 >> RS_DENSE_SHIFT;
      const bool split = b1 != b2;
      const rs_dblk k1 = rs_load_dense(ix, b1);
      rs_dblk k2 = k1;
      if (split) k2 = rs_load_dense(ix, b2); /* issued before anything waits on k1: two blocks in flight */
      const uint32_t t1 = ((uint32_t)p1 & (RS_DENSE_SYMS - 1u)) + 1u, t2 = ((uint32_t)p2 & (RS_DENSE_SYMS - 1u)) + 1u;
      uint32_t a1, a2;
      rs_dense_count2(k1, c, t1, split ? 0u : t2, &a1, &a2);
      uint32_t before2 = c == 1 ? k1.cnt[0] : c == 2 ? k1.cnt[1] : c == 3 ? k1.cnt[2] : k1.cnt[3];
      const P r1 = none1 ? 0 : (P)before2 + a1;
      if (split) { /* rare once the interval is narrow: lower-1 and upper then sit in one block */
        uint32_t unused;
        rs_dense_count2(k2, c, t2, 0u, &a2, &unused);
        before2 = c == 1 ? k2.cnt[0] : c == 2 ? k2.cnt[1] : c == 3 ? k2.cnt[2] : k2.cnt[3];
        ++nsplit;
      }
      const P r2 = none2 ? 0 : (P)before2 + a2;
      ++nsteps;
      const P pb = (P)ix.pc[c];
      lo = pb + r1;
      hi = pb + r2 - 1;
      --j;
      /* query.cpp:35-37 compares unsigned: an upper bound of 2^64-1 ("before row 0", NONE here) is above every lower
       * bound, so the reference searches on -- and so does this */
      if (lo > hi || j < 0) {
        rs_emit(mode, lower, upper, peers, q, (uint64_t)lo, hi == NONE ? ~0ull : (uint64_t)hi, pk, k, rc);
        active = false;
      }
    }
  }
  if (stats && nsteps) {
    atomicAdd(stats, (unsigned long long)nsteps);
    atomicAdd(stats + 2, (unsigned long long)(nsteps + nsplit)); /* 64-byte blocks read */
  }
}

/* One counter per launch, from a pool that is cleared in bulk: rs_ticket_take() hands out the next `n` consecutive
 * slots; a slot is reused only after RS_TICKET_POOL further launches, by which time its kernel has long ended (the
 * driver queues at most a few thousand launches). */
#define RS_TICKET_POOL 65536u
static unsigned long long* g_tickets[64] = {};
static uint64_t g_ticket_next[64] = {};
static std::mutex g_ticket_mu;

cudaError_t rs_ticket_take(uint32_t n, unsigned long long** first, cudaStream_t st) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64 || n == 0 || n > RS_TICKET_POOL / 4) return cudaErrorInvalidValue;
  std::lock_guard<std::mutex> lk(g_ticket_mu);
  if (!g_tickets[dev]) {
    e = cudaMalloc((void**)&g_tickets[dev], RS_TICKET_POOL * sizeof(unsigned long long));
    if (e != cudaSuccess) return e;
  }
  uint64_t at = g_ticket_next[dev] % RS_TICKET_POOL;
  if (at + n > RS_TICKET_POOL) at = 0; /* keep a request contiguous */
  g_ticket_next[dev] = at + n;
  *first = g_tickets[dev] + at;
  return cudaMemsetAsync(*first, 0, (size_t)n * sizeof(unsigned long long), st);
}

/* K1b.  The rare queries K1 left flagged (bit 63 of lower[q]: a byte outside "$ACGT"), one whole findInterval
 * each, as the reference answers them (rs_find_interval_one).  A kernel of its own so that this path costs K1's
 * loop no registers; when nothing is flagged it reads lower[] once (8 bytes per query, still in L2) and ends. */
__global__ void __launch_bounds__(RS_THREADS)
    k_search_flagged(const __grid_constant__ rs_index_dev ix, const uint8_t* __restrict__ text,
                     const uint64_t* __restrict__ offs, uint32_t k, uint64_t Q, uint64_t* __restrict__ lower,
                     uint64_t* __restrict__ upper, unsigned long long* __restrict__ stats) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  uint32_t nsteps = 0;
  for (uint64_t q = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; q < Q; q += stride) {
    if (!(lower[q] >> 63)) continue; /* a valid lower bound never has bit 63 set */
    const uint64_t base = offs ? offs[q] : q * (uint64_t)k;
    uint64_t lo, hi;
    nsteps += rs_find_interval_one(ix, text + base, offs ? offs[q + 1] - base : (uint64_t)k, &lo, &hi);
    lower[q] = lo;
    upper[q] = hi;
  }
  if (stats && nsteps) atomicAdd(stats, (unsigned long long)nsteps);
}

/* Builds the locate table (rs_layout.h): one walker per read, from the row of its terminator back to its first
 * base, leaving at every row it passes its own start row and the distance walked.  n LF steps in total. */
#define RS_LOC_STEP_MASK 0xFFFFFFull
#define RS_K2_BLOCKS_BUILD 4
__global__ void __launch_bounds__(RS_THREADS, RS_K2_BLOCKS_BUILD)
    k_build_locate(const __grid_constant__ rs_index_dev ix, uint32_t max_len, uint64_t* __restrict__ loc,
                   uint64_t* __restrict__ fin) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  uint64_t next_i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x;
  uint64_t i = 0, idx = 0;
  uint32_t s = 0;
  bool active = false;
  while (true) {
    if (!active && next_i < ix.num_strings) {
      i = next_i;
      next_i += stride;
      idx = i;
      s = 0;
      active = true;
    }
    if (__all_sync(RS_FULL, !active)) break;
    if (active) {
      /* under the reference's tail behaviour an LF step can land on row n: nothing to record there */
      if (idx < ix.n) loc[idx] = (i << 24) | s;
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      idx = nxt;
      if (sym == 0) {
        fin[i] = (idx << 24) | s;
        active = false;
      } else if (++s > max_len) { /* longer than any read: a walk caught in the reference's tail region */
        fin[i] = ~0ull;
        active = false;
      }
    }
  }
}

/* K2 with the locate table: one 8-byte read (+ one of loc_fin, + .sai) per row instead of a walk; rows the table
 * does not cover (walks that never end, under the reference's tail behaviour) are walked as K2 does. */
__global__ void __launch_bounds__(RS_THREADS)
    k_locate_lookup(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ rows, uint64_t n,
                    uint32_t max_steps, uint64_t* __restrict__ lexrank, uint32_t* __restrict__ steps,
                    const uint64_t* __restrict__ sai, uint64_t* __restrict__ read_index,
                    unsigned long long* __restrict__ stats, const uint64_t* __restrict__ n_dev) {
  if (n_dev && *n_dev < n) n = *n_dev; /* item count left on the device by an earlier kernel (rs_pop.cu) */
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  unsigned long long lf_steps = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
    uint64_t idx = rows[i];
    uint64_t term = ~0ull;
    uint32_t nsteps = 0xFFFFFFFFu;
    bool known = false;
    if (idx < ix.n) {
      const uint64_t e = __ldg(ix.loc + idx);
      if (e != ~0ull) {
        const uint64_t f = __ldg(ix.loc_fin + (e >> 24));
        if (f != ~0ull) {
          known = true;
          const uint32_t left = (uint32_t)(f & RS_LOC_STEP_MASK) - (uint32_t)(e & RS_LOC_STEP_MASK);
          if (left <= max_steps) {
            term = f >> 24;
            nsteps = left;
          }
        }
      }
    }
    if (!known) { /* extractPrefix as K2 walks it */
      uint32_t s = 0;
      while (true) {
        uint64_t nxt;
        const uint32_t sym = rs_lf_step(ix, idx, &nxt);
        ++lf_steps;
        idx = nxt;
        if (sym == 0) {
          term = idx;
          nsteps = s;
          break;
        }
        if (++s > max_steps) break;
      }
    }
    if (lexrank) lexrank[i] = term;
    if (steps) steps[i] = nsteps;
    if (read_index) read_index[i] = (term != ~0ull && term < ix.num_strings) ? sai[term] : ~0ull;
  }
  if (stats && lf_steps) atomicAdd(stats + 1, lf_steps);
}

/* ---- sampled locate table ---- */
/* One 16-byte record per 64 rows: {marker bits, sampled rows before the word}.  A walk step fetches the record of its
 * row together with the row's rank block; when the row is sampled the record alone gives the slot of its entry
 * (round 1 kept the bits and a popcount directory per 512 rows apart: up to eight more words on every sampled row). */
/* pass 1: mark the rows every `every` symbols along each read's walk; fin as in the full table */
__global__ void __launch_bounds__(RS_THREADS, RS_K2_BLOCKS_BUILD)
    k_mark_locate(const __grid_constant__ rs_index_dev ix, uint32_t max_len, uint32_t every,
                  ulonglong2* __restrict__ rec, uint64_t* __restrict__ fin) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < ix.num_strings; i += stride) {
    uint64_t idx = i;
    uint32_t s = 0;
    while (true) {
      if (s % every == 0 && idx < ix.n)
        atomicOr(reinterpret_cast<unsigned long long*>(&rec[idx >> 6].x), 1ull << (idx & 63));
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      idx = nxt;
      if (sym == 0) {
        fin[i] = (idx << 24) | s;
        break;
      }
      if (++s > max_len) {
        fin[i] = ~0ull;
        break;
      }
    }
  }
}

/* set bits of each word (input of the scan that yields the records' second halves) */
__global__ void k_mark_counts(const ulonglong2* __restrict__ rec, uint64_t nwords, uint64_t* __restrict__ counts) {
  const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < nwords) counts[w] = __popcll(rec[w].x);
}
__global__ void k_mark_ranks(ulonglong2* __restrict__ rec, uint64_t nwords, const uint64_t* __restrict__ before) {
  const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < nwords) rec[w].y = before[w];
}

__device__ __forceinline__ uint64_t rs_sample_slot(const ulonglong2 r, uint64_t row) {
  return r.y + __popcll(r.x & ((1ull << (row & 63)) - 1));
}

/* pass 2: the same walks again, filling the entries of the sampled rows */
__global__ void __launch_bounds__(RS_THREADS, RS_K2_BLOCKS_BUILD)
    k_fill_locate(const __grid_constant__ rs_index_dev ix, uint32_t max_len, uint32_t every,
                  const ulonglong2* __restrict__ rec, uint64_t* __restrict__ loc) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < ix.num_strings; i += stride) {
    uint64_t idx = i;
    uint32_t s = 0;
    while (true) {
      if (s % every == 0 && idx < ix.n) loc[rs_sample_slot(rec[idx >> 6], idx)] = (i << 24) | s;
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      idx = nxt;
      if (sym == 0 || ++s > max_len) break;
    }
  }
}

/* K2 with the sampled table: follow the walk until a sampled row (or the '$').  Rows are handed out from a counter
 * as in K1 (a warp takes as many as it has idle lanes), so the lanes stay busy although walks end after 0 .. every - 1
 * steps. */
__global__ void __launch_bounds__(RS_THREADS, RS_K2_BLOCKS_BUILD)
    k_locate_sampled(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ rows, uint64_t n,
                     uint32_t max_steps, uint64_t* __restrict__ lexrank, uint32_t* __restrict__ steps,
                     const uint64_t* __restrict__ sai, uint64_t* __restrict__ read_index,
                     unsigned long long* __restrict__ stats, const uint64_t* __restrict__ n_dev,
                     unsigned long long* __restrict__ ticket) {
  if (n_dev && *n_dev < n) n = *n_dev; /* item count left on the device by an earlier kernel (rs_pop.cu) */
  const ulonglong2* __restrict__ rec = reinterpret_cast<const ulonglong2*>(ix.loc_mark);
  const uint32_t lane = threadIdx.x & 31u;
  unsigned long long lf_steps = 0;
  uint64_t i = 0, idx = 0;
  uint32_t s = 0;
  bool active = false, exhausted = false;
  while (true) {
    const uint32_t idle = __ballot_sync(RS_FULL, !active);
    if (idle == RS_FULL || __popc(idle) >= RS_REFILL_MIN) {
      if (!exhausted) {
        const int leader = __ffs((int)idle) - 1;
        unsigned long long base = 0;
        if ((int)lane == leader) base = atomicAdd(ticket, (unsigned long long)__popc(idle));
        base = __shfl_sync(RS_FULL, base, leader);
        if (base >= n) exhausted = true;
        if (!active && !exhausted) {
          i = base + (uint64_t)__popc(idle & ((1u << lane) - 1u));
          if (i < n) {
            idx = rows[i];
            s = 0;
            active = true;
          }
        }
      }
      if (__all_sync(RS_FULL, !active)) {
        if (exhausted) break;
        continue;
      }
    }
    if (active) {
      /* the record and the block of the LF step are fetched together: the step is wasted on a sampled row, but the
       * two reads do not wait for each other on the others */
      const ulonglong2 r = idx < ix.n ? __ldg(rec + (idx >> 6)) : make_ulonglong2(0, 0);
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      bool done = false;
      uint64_t term = ~0ull;
      uint32_t nsteps = 0xFFFFFFFFu;
      if ((r.x >> (idx & 63)) & 1ull) {
        const uint64_t e = __ldg(ix.loc + rs_sample_slot(r, idx));
        const uint64_t f = e != ~0ull ? __ldg(ix.loc_fin + (e >> 24)) : ~0ull;
        if (f != ~0ull) {
          const uint64_t total = (uint64_t)s + ((f & RS_LOC_STEP_MASK) - (e & RS_LOC_STEP_MASK));
          if (total <= max_steps) {
            term = f >> 24;
            nsteps = (uint32_t)total;
          }
          done = true;
        }
      }
      if (!done) {
        ++lf_steps;
        idx = nxt;
        if (sym == 0) {
          term = idx;
          nsteps = s;
          done = true;
        } else if (++s > max_steps) {
          done = true;
        }
      }
      if (done) {
        if (lexrank) lexrank[i] = term;
        if (steps) steps[i] = nsteps;
        if (read_index) read_index[i] = (term != ~0ull && term < ix.num_strings) ? sai[term] : ~0ull;
        active = false;
      }
    }
  }
  if (stats && lf_steps) atomicAdd(stats + 1, lf_steps);
}

/* K2.  extractPrefix (query.cpp:43-63): LF-walk from each row to its '$'.  Same step-interleaved
 * scheme as K1: a thread whose walk ends takes its next row on the following trip, so the set of
 * live walkers stays dense in every warp without a separate compaction pass. */
#ifndef RS_K2_BLOCKS
#define RS_K2_BLOCKS 5 /* 4: 1119, 5: 1164, 6: 1012 M hit rows/s */
#endif
__global__ void __launch_bounds__(RS_THREADS, RS_K2_BLOCKS)
    k_extract_prefix(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ rows, uint64_t n,
                     uint32_t max_steps, uint64_t* __restrict__ lexrank, uint32_t* __restrict__ steps,
                     const uint64_t* __restrict__ sai, uint64_t* __restrict__ read_index, uint8_t* __restrict__ bases,
                     unsigned long long* __restrict__ stats, const uint64_t* __restrict__ n_dev,
                     unsigned long long* __restrict__ ticket) {
  if (n_dev && *n_dev < n) n = *n_dev;
  const uint32_t lane = threadIdx.x & 31u;
  unsigned long long lf_steps = 0;
  uint64_t i = 0, idx = 0;
  uint32_t s = 0;
  bool active = false, exhausted = false;
  while (true) {
    /* rows are handed out from one counter, a warp taking as many as it has idle lanes (walks are 1 .. read length
     * steps long: with a fixed share of the rows per thread a warp ran until its slowest lane was done) */
    const uint32_t idle = __ballot_sync(RS_FULL, !active);
    if (idle == RS_FULL || __popc(idle) >= RS_REFILL_MIN) {
      if (!exhausted) {
        const int leader = __ffs((int)idle) - 1;
        unsigned long long base = 0;
        if ((int)lane == leader) base = atomicAdd(ticket, (unsigned long long)__popc(idle));
        base = __shfl_sync(RS_FULL, base, leader);
        if (base >= n) exhausted = true;
        if (!active && !exhausted) {
          i = base + (uint64_t)__popc(idle & ((1u << lane) - 1u));
          if (i < n) {
            idx = rows[i];
            s = 0;
            active = true;
          }
        }
      }
      if (__all_sync(RS_FULL, !active)) {
        if (exhausted) break;
        continue;
      }
    }
    if (active) {
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      ++lf_steps;
      idx = nxt;
      bool done = false;
      uint64_t term = 0;
      uint32_t nsteps = 0;
      if (sym == 0) {
        done = true;
        term = idx;
        nsteps = s;
      } else {
        if (bases && s < max_steps) bases[i * (uint64_t)max_steps + s] = (uint8_t)rs_ascii(sym);
        ++s;
        if (s > max_steps) { /* the reference would never return (walk cycling in its tail region) */
          done = true;
          term = ~0ull;
          nsteps = 0xFFFFFFFFu;
        }
      }
      if (done) {
        if (lexrank) lexrank[i] = term;
        if (steps) steps[i] = nsteps;
        if (read_index) read_index[i] = (term != ~0ull && term < ix.num_strings) ? sai[term] : ~0ull;
        if (bases) { /* std::reverse (query.cpp:60): the walk produced the prefix back to front */
          uint8_t* p = bases + i * (uint64_t)max_steps;
          const uint32_t m = nsteps == 0xFFFFFFFFu ? 0u : nsteps;
          for (uint32_t a = 0; a < m / 2; ++a) {
            const uint8_t t = p[a];
            p[a] = p[m - 1 - a];
            p[m - 1 - a] = t;
          }
          for (uint32_t a = m; a < max_steps; ++a) p[a] = 0;
        }
        active = false;
      }
    }
  }
  if (stats && lf_steps) atomicAdd(stats + 1, lf_steps);
}

/* extractPostfix (query.cpp:65-85): forward walk by getF / getOccAt from each row to the '$' row of
 * its read.  read_index = the row at which getF returned '$' (the read's index in builder input
 * order); the select of that last step is still performed, as the reference does. */
__global__ void __launch_bounds__(RS_THREADS)
    k_extract_postfix(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ rows, uint64_t n,
                      uint32_t max_steps, uint64_t* __restrict__ read_index, uint32_t* __restrict__ steps,
                      uint8_t* __restrict__ bases) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  uint64_t next_i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x;
  uint64_t i = 0, idx = 0;
  uint32_t s = 0;
  bool active = false;
  while (true) {
    if (!active && next_i < n) {
      i = next_i;
      next_i += stride;
      idx = rows[i];
      s = 0;
      active = true;
    }
    if (__all_sync(RS_FULL, !active)) break;
    if (active) {
      uint64_t nxt;
      const uint64_t at = idx;
      const uint32_t f = rs_fl_step(ix, idx < ix.n ? idx : ix.n - 1, &nxt);
      idx = nxt;
      bool done = false;
      uint64_t term = 0;
      uint32_t nsteps = 0;
      if (f == 0) {
        done = true;
        term = at;
        nsteps = s;
      } else {
        if (bases && s < max_steps) bases[i * (uint64_t)max_steps + s] = (uint8_t)rs_ascii(f);
        ++s;
        if (s > max_steps || idx == ~0ull) {
          done = true;
          term = ~0ull;
          nsteps = 0xFFFFFFFFu;
        }
      }
      if (done) {
        if (read_index) read_index[i] = term;
        if (steps) steps[i] = nsteps;
        if (bases) {
          uint8_t* p = bases + i * (uint64_t)max_steps;
          const uint32_t m = nsteps == 0xFFFFFFFFu ? 0u : nsteps;
          for (uint32_t a = m; a < max_steps; ++a) p[a] = 0;
        }
        active = false;
      }
    }
  }
}

/* ---- launchers ---- */
static int rs_grid_for(const void* kernel, uint64_t items) {
  /* SM count and per-kernel occupancy are looked up once (the GPUs of a box are identical) */
  static int sms = 0;
  static struct { const void* k; int per_sm; } cache[16];
  static int ncache = 0;
  static std::mutex mu;
  int per_sm = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!sms) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (sms <= 0) sms = 148;
    }
    for (int i = 0; i < ncache; ++i)
      if (cache[i].k == kernel) per_sm = cache[i].per_sm;
    if (!per_sm) {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, RS_THREADS, 0);
      if (per_sm < 1) per_sm = 1;
      if (ncache < 16) cache[ncache++] = {kernel, per_sm};
    }
  }
  uint64_t grid = (uint64_t)sms * per_sm; /* one resident wave, a multiple of the SM count */
  const uint64_t need = (items + RS_THREADS - 1) / RS_THREADS;
  if (need < grid) grid = need;
  return grid ? (int)grid : 1;
}

#define RS_LAUNCH(KERNEL, ITEMS, ST, ...) \
  KERNEL<<<rs_grid_for((const void*)KERNEL, ITEMS), RS_THREADS, 0, ST>>>(__VA_ARGS__)

cudaError_t rs_launch_occ(const rs_index_dev& ix, const uint8_t* syms, const uint64_t* idx, uint64_t n, uint64_t* out,
                          cudaStream_t st) {
  if (!n) return cudaSuccess;
  RS_LAUNCH(k_occ, n, st, ix, syms, idx, n, out);
  return cudaGetLastError();
}

cudaError_t rs_launch_char(const rs_index_dev& ix, const uint64_t* idx, uint64_t n, uint8_t* out, cudaStream_t st) {
  if (!n) return cudaSuccess;
  RS_LAUNCH(k_char, n, st, ix, idx, n, out);
  return cudaGetLastError();
}

cudaError_t rs_launch_getf(const rs_index_dev& ix, const uint64_t* idx, uint64_t n, uint8_t* out, cudaStream_t st) {
  if (!n) return cudaSuccess;
  k_getf<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ix, idx, n, out);
  return cudaGetLastError();
}

cudaError_t rs_launch_find_interval(const rs_index_dev& ix, const uint8_t* text, const uint64_t* offs, uint32_t k,
                                    uint64_t Q, uint64_t* lower, uint64_t* upper, int mode, int strand,
                                    unsigned long long* stats, cudaStream_t st, const rs_peers* peers_in) {
  if (!Q) return cudaSuccess;
  rs_peers peers;
  memset(&peers, 0, sizeof(peers));
  if (peers_in) peers = *peers_in;
  const bool packed_already = (mode == 3); /* mode 3 = mode 2 over keys an earlier call left in lower[] */
  if (packed_already) mode = 2;
  if (mode == 2 && (!peers_in || offs || k > RS_PACK_MAX_K)) return cudaErrorInvalidValue;
  if (mode == 4) return cudaErrorInvalidValue; /* rs_launch_find_interval_packed */
  const bool ftab = ix.ftab && ix.ftab_k > 1;
  if (!offs && k >= 1 && k <= RS_PACK_MAX_K) {
    if (!packed_already)
      k_pack_kmers<<<(unsigned)((Q + RS_THREADS - 1) / RS_THREADS), RS_THREADS, 0, st>>>(text, k, Q, mode >= 1 && strand == 1,
                                                                                     lower);
    if (ix.dense) {
      unsigned long long* ticket = nullptr;
      const cudaError_t te = rs_ticket_take(1, &ticket, st);
      if (te != cudaSuccess) return te;
      const bool small = ix.n < (1ull << 31);
      if (ftab && small) RS_LAUNCH((k_find_interval_dense<true, true>), Q, st, ix, k, Q, lower, lower, upper, mode, strand, stats, peers, ticket);
      else if (ftab) RS_LAUNCH((k_find_interval_dense<true, false>), Q, st, ix, k, Q, lower, lower, upper, mode, strand, stats, peers, ticket);
      else if (small) RS_LAUNCH((k_find_interval_dense<false, true>), Q, st, ix, k, Q, lower, lower, upper, mode, strand, stats, peers, ticket);
      else RS_LAUNCH((k_find_interval_dense<false, false>), Q, st, ix, k, Q, lower, lower, upper, mode, strand, stats, peers, ticket);
    } else if (ftab) RS_LAUNCH((k_find_interval<true, true>), Q, st, ix, text, offs, k, Q, lower, lower, upper, mode, strand, stats, peers);
    else RS_LAUNCH((k_find_interval<false, true>), Q, st, ix, text, offs, k, Q, lower, lower, upper, mode, strand, stats, peers);
  } else {
    if (ftab) RS_LAUNCH((k_find_interval<true, false>), Q, st, ix, text, offs, k, Q, lower, lower, upper, mode, strand, stats, peers);
    else RS_LAUNCH((k_find_interval<false, false>), Q, st, ix, text, offs, k, Q, lower, lower, upper, mode, strand, stats, peers);
  }
  if (mode == 0) {
    uint64_t blocks = (Q + 4 * RS_THREADS - 1) / (4 * RS_THREADS);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_search_flagged<<<(unsigned)blocks, RS_THREADS, 0, st>>>(ix, text, offs, k, Q, lower, upper, stats);
  }
  return cudaGetLastError();
}

cudaError_t rs_launch_pack_kmers(const uint8_t* text, uint32_t k, uint64_t Q, int rc, uint64_t* packed, cudaStream_t st) {
  if (!Q) return cudaSuccess;
  if (k < 1 || k > RS_PACK_MAX_K) return cudaErrorInvalidValue;
  k_pack_kmers<<<(unsigned)((Q + RS_THREADS - 1) / RS_THREADS), RS_THREADS, 0, st>>>(text, k, Q, rc, packed);
  return cudaGetLastError();
}

/* K1 over packed keys kept apart from the outputs (one batch searched in many shards: rs_pop.cu).
 * mode 4: first[q] = lower bound, rows[q] = rows of the interval (0 for a key flagged as holding a byte outside
 * ACGT: count_reads, service.cpp:299-301).  mode 2: counts added into the owners' buffers (peers). */
cudaError_t rs_launch_find_interval_packed(const rs_index_dev& ix, const uint64_t* pkeys, uint32_t k, uint64_t Q,
                                           uint64_t* first, uint32_t* rows, int mode, unsigned long long* stats,
                                           cudaStream_t st, const rs_peers* peers_in, unsigned long long* ticket) {
  if (!Q) return cudaSuccess;
  if (k < 1 || k > RS_PACK_MAX_K || (mode != 4 && mode != 2) || (mode == 2 && !peers_in)) return cudaErrorInvalidValue;
  rs_peers peers;
  memset(&peers, 0, sizeof(peers));
  if (peers_in) peers = *peers_in;
  const bool ftab = ix.ftab && ix.ftab_k > 1;
  uint64_t* up = reinterpret_cast<uint64_t*>(rows);
  if (ix.dense) {
    if (!ticket) { /* the caller did not reserve (and clear) one with its other launches */
      const cudaError_t te = rs_ticket_take(1, &ticket, st);
      if (te != cudaSuccess) return te;
    }
    const bool small = ix.n < (1ull << 31);
    if (ftab && small) RS_LAUNCH((k_find_interval_dense<true, true>), Q, st, ix, k, Q, pkeys, first, up, mode, 0, stats, peers, ticket);
    else if (ftab) RS_LAUNCH((k_find_interval_dense<true, false>), Q, st, ix, k, Q, pkeys, first, up, mode, 0, stats, peers, ticket);
    else if (small) RS_LAUNCH((k_find_interval_dense<false, true>), Q, st, ix, k, Q, pkeys, first, up, mode, 0, stats, peers, ticket);
    else RS_LAUNCH((k_find_interval_dense<false, false>), Q, st, ix, k, Q, pkeys, first, up, mode, 0, stats, peers, ticket);
  } else if (ftab) RS_LAUNCH((k_find_interval<true, true>), Q, st, ix, nullptr, nullptr, k, Q, pkeys, first, up, mode, 0, stats, peers);
  else RS_LAUNCH((k_find_interval<false, true>), Q, st, ix, nullptr, nullptr, k, Q, pkeys, first, up, mode, 0, stats, peers);
  return cudaGetLastError();
}

/* all 4^t strings of length t over ACGT, string number i at out + i*t (first character most significant) */
__global__ void k_all_kmers(uint32_t t, uint64_t n, uint8_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t v = i;
  for (int a = (int)t - 1; a >= 0; --a) {
    out[i * t + a] = (uint8_t)rs_ascii(1u + (uint32_t)(v & 3u));
    v >>= 2;
  }
}

__global__ void k_interleave(const uint64_t* __restrict__ lo, const uint64_t* __restrict__ up, uint64_t n,
                             ulonglong2* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = make_ulonglong2(lo[i], up[i]);
}

/* Fills table[4^t] with findInterval of every t-mer, computed by K1 itself WITHOUT a table (ix.ftab is ignored). */
cudaError_t rs_build_ftab(const rs_index_dev& ix_in, uint32_t t, uint64_t* table, cudaStream_t st) {
  rs_index_dev ix = ix_in;
  ix.ftab = nullptr;
  ix.ftab_k = 0;
  const uint64_t n = 1ull << (2 * t);
  uint8_t* kmers = nullptr;
  uint64_t *lo = nullptr, *up = nullptr;
  cudaError_t e;
  e = cudaMallocAsync((void**)&kmers, n * t, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void**)&lo, n * 8, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void**)&up, n * 8, st);
  if (e != cudaSuccess) {
    if (kmers) cudaFreeAsync(kmers, st);
    if (lo) cudaFreeAsync(lo, st);
    return e;
  }
  k_all_kmers<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, n, kmers);
  e = rs_launch_find_interval(ix, kmers, nullptr, t, n, lo, up, 0, 0, nullptr, st, nullptr);
  if (e == cudaSuccess) {
    k_interleave<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(lo, up, n, reinterpret_cast<ulonglong2*>(table));
    e = cudaGetLastError();
  }
  cudaFreeAsync(kmers, st);
  cudaFreeAsync(lo, st);
  cudaFreeAsync(up, st);
  return e;
}

/* n_dev (nullable): device word holding the number of rows (<= n, which then only bounds the grid) */
cudaError_t rs_launch_extract_prefix_n(const rs_index_dev& ix, const uint64_t* rows, const uint64_t* n_dev, uint64_t n,
                                       uint32_t max_steps, const uint64_t* sai, uint64_t* read_index,
                                       unsigned long long* stats, cudaStream_t st) {
  if (!n) return cudaSuccess;
  unsigned long long* ticket = nullptr;
  if (!(ix.loc && ix.loc_every == 1)) {
    const cudaError_t te = rs_ticket_take(1, &ticket, st);
    if (te != cudaSuccess) return te;
  }
  if (ix.loc && ix.loc_every > 1) {
    RS_LAUNCH(k_locate_sampled, n, st, ix, rows, n, max_steps, nullptr, nullptr, sai, read_index, stats, n_dev, ticket);
  } else if (ix.loc) {
    uint64_t blocks = (n + RS_THREADS - 1) / RS_THREADS;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_locate_lookup<<<(unsigned)blocks, RS_THREADS, 0, st>>>(ix, rows, n, max_steps, nullptr, nullptr, sai, read_index, stats, n_dev);
  } else {
    RS_LAUNCH(k_extract_prefix, n, st, ix, rows, n, max_steps, nullptr, nullptr, sai, read_index, nullptr, stats, n_dev, ticket);
  }
  return cudaGetLastError();
}

cudaError_t rs_launch_extract_prefix(const rs_index_dev& ix, const uint64_t* rows, uint64_t n, uint32_t max_steps,
                                     uint64_t* lexrank, uint32_t* steps, const uint64_t* sai, uint64_t* read_index,
                                     uint8_t* bases, unsigned long long* stats, cudaStream_t st) {
  if (!n) return cudaSuccess;
  const uint64_t* none = nullptr;
  unsigned long long* ticket = nullptr;
  if (!(ix.loc && ix.loc_every == 1 && !bases)) {
    const cudaError_t te = rs_ticket_take(1, &ticket, st);
    if (te != cudaSuccess) return te;
  }
  if (ix.loc && ix.loc_every > 1 && !bases) {
    RS_LAUNCH(k_locate_sampled, n, st, ix, rows, n, max_steps, lexrank, steps, sai, read_index, stats, none, ticket);
    return cudaGetLastError();
  }
  if (ix.loc && !bases) {
    uint64_t blocks = (n + RS_THREADS - 1) / RS_THREADS;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_locate_lookup<<<(unsigned)blocks, RS_THREADS, 0, st>>>(ix, rows, n, max_steps, lexrank, steps, sai, read_index, stats, none);
    return cudaGetLastError();
  }
  RS_LAUNCH(k_extract_prefix, n, st, ix, rows, n, max_steps, lexrank, steps, sai, read_index, bases, stats, none, ticket);
  return cudaGetLastError();
}

/* sampled table, pass 1: records (zeroed here) and fin[]; the caller sizes loc from the marks */
cudaError_t rs_locate_mark(const rs_index_dev& ix_in, uint32_t max_len, uint32_t every, uint64_t* rec, uint64_t* fin,
                           cudaStream_t st) {
  rs_index_dev ix = ix_in;
  ix.loc = nullptr;
  ix.loc_every = 0;
  cudaError_t e;
  const uint64_t nwords = (ix.n + 63) / 64;
  if ((e = cudaMemsetAsync(rec, 0, nwords * 16, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(fin, 0xFF, (ix.num_strings ? ix.num_strings : 1) * 8, st)) != cudaSuccess) return e;
  if (!ix.num_strings) return cudaSuccess;
  RS_LAUNCH(k_mark_locate, ix.num_strings, st, ix, max_len, every, reinterpret_cast<ulonglong2*>(rec), fin);
  return cudaGetLastError();
}

/* counts[w] = marked rows of word w; the caller scans them and hands the result to rs_locate_mark_ranks */
cudaError_t rs_locate_mark_counts(const uint64_t* rec, uint64_t n, uint64_t* counts, cudaStream_t st) {
  const uint64_t nwords = (n + 63) / 64;
  k_mark_counts<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(reinterpret_cast<const ulonglong2*>(rec), nwords, counts);
  return cudaGetLastError();
}
cudaError_t rs_locate_mark_ranks(uint64_t* rec, uint64_t n, const uint64_t* before, cudaStream_t st) {
  const uint64_t nwords = (n + 63) / 64;
  k_mark_ranks<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(reinterpret_cast<ulonglong2*>(rec), nwords, before);
  return cudaGetLastError();
}

/* sampled table, pass 2 */
cudaError_t rs_locate_fill(const rs_index_dev& ix_in, uint32_t max_len, uint32_t every, const uint64_t* rec,
                           uint64_t* loc, uint64_t nsamples, cudaStream_t st) {
  rs_index_dev ix = ix_in;
  ix.loc = nullptr;
  ix.loc_every = 0;
  cudaError_t e;
  if ((e = cudaMemsetAsync(loc, 0xFF, (nsamples ? nsamples : 1) * 8, st)) != cudaSuccess) return e;
  if (!ix.num_strings) return cudaSuccess;
  RS_LAUNCH(k_fill_locate, ix.num_strings, st, ix, max_len, every, reinterpret_cast<const ulonglong2*>(rec), loc);
  return cudaGetLastError();
}

cudaError_t rs_build_locate_table(const rs_index_dev& ix_in, uint32_t max_len, uint64_t* loc, uint64_t* fin, cudaStream_t st) {
  rs_index_dev ix = ix_in;
  ix.loc = nullptr;
  ix.loc_fin = nullptr;
  ix.loc_every = 0;
  cudaError_t e;
  if ((e = cudaMemsetAsync(loc, 0xFF, ix.n * 8, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(fin, 0xFF, (ix.num_strings ? ix.num_strings : 1) * 8, st)) != cudaSuccess) return e;
  if (!ix.num_strings) return cudaSuccess;
  RS_LAUNCH(k_build_locate, ix.num_strings, st, ix, max_len, loc, fin);
  return cudaGetLastError();
}

cudaError_t rs_launch_occ_at(const rs_index_dev& ix, const uint8_t* syms, const uint64_t* cnt, uint64_t n, uint64_t* out,
                             cudaStream_t st) {
  if (!n) return cudaSuccess;
  RS_LAUNCH(k_occ_at, n, st, ix, syms, cnt, n, out);
  return cudaGetLastError();
}

cudaError_t rs_launch_extract_postfix(const rs_index_dev& ix, const uint64_t* rows, uint64_t n, uint32_t max_steps,
                                      uint64_t* read_index, uint32_t* steps, uint8_t* bases, cudaStream_t st) {
  if (!n) return cudaSuccess;
  RS_LAUNCH(k_extract_postfix, n, st, ix, rows, n, max_steps, read_index, steps, bases);
  return cudaGetLastError();
}

/* Dense rank blocks from the run blocks: one thread per dense block.  Header = exact ranks at the block's first
 * position (through the run blocks); planes = the 128 symbols, decoded run by run. */
__global__ void __launch_bounds__(RS_THREADS)
    k_dense_build(const __grid_constant__ rs_index_dev ix, uint64_t ndense, uint4* __restrict__ out, unsigned int* __restrict__ bad) {
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t b = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; b < ndense; b += stride) {
    const uint64_t P0 = b << RS_DENSE_SHIFT;
    uint32_t cnt[4] = {0, 0, 0, 0}, p0[4] = {0, 0, 0, 0}, p1[4] = {0, 0, 0, 0}, p2[4] = {0, 0, 0, 0};
    if (P0) {
      for (uint32_t c = 1; c <= 4; ++c) {
        const uint64_t r = rs_rank(ix, c, P0 - 1);
        if (r >> 32) atomicOr(bad, 1u);
        cnt[c - 1] = (uint32_t)r;
      }
    }
    uint64_t pos = P0;
    const uint64_t end = P0 + RS_DENSE_SYMS < ix.n ? P0 + RS_DENSE_SYMS : ix.n;
    uint64_t rb = rs_find_block(ix, pos);
    while (pos < end) {
      const uint8_t* p = rs_block_ptr(ix, rb);
      uint64_t at = rs_ld64(reinterpret_cast<const uint64_t*>(p)); /* first symbol of the run in hand */
      for (uint32_t t = 0; t < RS_PAYLOAD && pos < end; ++t) {
        const uint32_t v = p[RS_HDR_BYTES + t], len = v & 31u, sym = v >> 5;
        if (!len) break; /* padding of the last block */
        const uint64_t stop = at + len;
        while (pos < stop && pos < end) { /* pos >= at: the block was found by position */
          const uint32_t i = (uint32_t)(pos - P0), w = i >> 5, bit = 1u << (i & 31u);
          if (sym == 0) p2[w] |= bit;
          else {
            if ((sym - 1u) & 1u) p0[w] |= bit;
            if ((sym - 1u) & 2u) p1[w] |= bit;
          }
          ++pos;
        }
        at = stop;
      }
      ++rb;
    }
    out[4 * b + 0] = make_uint4(cnt[0], cnt[1], cnt[2], cnt[3]);
    out[4 * b + 1] = make_uint4(p0[0], p0[1], p0[2], p0[3]);
    out[4 * b + 2] = make_uint4(p1[0], p1[1], p1[2], p1[3]);
    out[4 * b + 3] = make_uint4(p2[0], p2[1], p2[2], p2[3]);
  }
}

cudaError_t rs_build_dense(const rs_index_dev& ix_in, uint64_t ndense, void* out, unsigned int* bad, cudaStream_t st) {
  rs_index_dev ix = ix_in;
  ix.dense = nullptr; /* built FROM the run blocks */
  if (!ndense) return cudaSuccess;
  RS_LAUNCH(k_dense_build, ndense, st, ix, ndense, reinterpret_cast<uint4*>(out), bad);
  return cudaGetLastError();
}

/* ---- count without any allocation (rs_count_batch_noalloc_dev) ----
 * The count path's scratch -- the packed keys K1 searches (8 bytes per query) and the work counter the dense K1 hands
 * its queries out from -- lives in a file-scope device array instead of cudaMalloc'd memory: callers that may not
 * allocate (under stream capture, in a serving loop) can still count.  Every GPU that loads the library holds the
 * array (8 MiB + 256 B) until the process ends.  ALL calls on a device share it: two calls that overlap on the GPU
 * (different streams, no ordering between them) race on the keys and on the counter; calls on one stream, or ordered
 * by events, are safe.  The array keeps the previous call's contents: the counter is cleared by every call, the keys
 * of queries 0..Q-1 are rewritten by every call before K1 reads them.
 * `g_noalloc_scratch` is a host-side stand-in in host code, not a device address: the address is taken ONCE per
 * device with cudaGetSymbolAddress, with that device current. */
#define RS_NOALLOC_HEAD 256u /* the work counter (first 8 bytes), keeping the keys 256-byte aligned */
static __device__ __align__(256) unsigned char g_noalloc_scratch[RS_NOALLOC_HEAD + (size_t)RS_NOALLOC_MAX_QUERIES * 8u];
static unsigned char* g_noalloc_ptr[64] = {};
static std::mutex g_noalloc_mu;

cudaError_t rs_launch_count_noalloc(const rs_index_dev& ix, const uint8_t* text, uint32_t k, uint64_t Q, int strand,
                                    uint64_t* counts, unsigned long long* stats, cudaStream_t st) {
  if (!Q) return cudaSuccess;
  if (k < 1 || k > RS_PACK_MAX_K || Q > RS_NOALLOC_MAX_QUERIES || (strand != 0 && strand != 1)) return cudaErrorInvalidValue;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  unsigned char* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) { /* set-up, once per device, that device being current */
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e != cudaSuccess) return e;
      g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    scratch = g_noalloc_ptr[dev];
  }
  unsigned long long* ticket = reinterpret_cast<unsigned long long*>(scratch);
  uint64_t* pkeys = reinterpret_cast<uint64_t*>(scratch + RS_NOALLOC_HEAD);
  e = cudaMemsetAsync(ticket, 0, RS_NOALLOC_HEAD, st); /* the counter must start at zero; the last call left it at >= Q */
  if (e != cudaSuccess) return e;
  k_pack_kmers<<<(unsigned)((Q + RS_THREADS - 1) / RS_THREADS), RS_THREADS, 0, st>>>(text, k, Q, strand == 1, pkeys);
  rs_peers peers;
  memset(&peers, 0, sizeof(peers));
  const bool ftab = ix.ftab && ix.ftab_k > 1;
  /* mode 1 (count) over keys kept apart from the output: counts[q] = rows of the interval, 0 for a flagged key */
  if (ix.dense) {
    const bool small = ix.n < (1ull << 31);
    if (ftab && small) RS_LAUNCH((k_find_interval_dense<true, true>), Q, st, ix, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers, ticket);
    else if (ftab) RS_LAUNCH((k_find_interval_dense<true, false>), Q, st, ix, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers, ticket);
    else if (small) RS_LAUNCH((k_find_interval_dense<false, true>), Q, st, ix, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers, ticket);
    else RS_LAUNCH((k_find_interval_dense<false, false>), Q, st, ix, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers, ticket);
  } else if (ftab) RS_LAUNCH((k_find_interval<true, true>), Q, st, ix, nullptr, nullptr, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers);
  else RS_LAUNCH((k_find_interval<false, true>), Q, st, ix, nullptr, nullptr, k, Q, pkeys, counts, nullptr, 1, strand, stats, peers);
  return cudaGetLastError();
}
