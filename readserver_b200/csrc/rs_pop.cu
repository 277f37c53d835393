/* Population queries over a SET of BWT shards resident on one GPU: count + locate + sample sets in one
 * device-side pipeline (include/readserver_b200.h, rs_shardset_* / rs_pop_*).
 *
 * The reference runs one worker process per shard; the front end broadcasts every query to every worker
 * (src/service/server.cpp:124,957) and, per query, sums the workers' counts and concatenates their read lists
 * (server.cpp:184-197, 404-410); each worker answers with findInterval (src/bwt/query.cpp:24-41), one walk per
 * hit row (extractPrefix / extractPostfix, query.cpp:43-85; find_reads, service.cpp:714-751) and one RocksDB Get +
 * value decode per read (service.cpp:816-845).  Here the shards of a box live in the HBM of its GPUs and one call
 * answers a whole batch:
 *
 *   k_pack_kmers          the batch becomes one u64 per k-mer, once
 *   per chunk of queries (the chunk's per-shard intervals stay in L2-sized scratch):
 *     K1 per shard        first row + row count of every query in that shard (mode 4; the shard's blocks, directory
 *                         and suffix table -- tens of MB -- stay L2-resident while the chunk streams through)
 *     k_pop_totals        population count = sum over shards; hits per query (rows capped at max_hits per shard)
 *     CUB scan            hit offsets (query-major: a query's hits are contiguous, shard after shard, row order)
 *     k_pop_gather        read index of every hit row and the union of the reads' sample sets:
 *                           read-index table (4 B per BWT row): an interval's rows are contiguous, so the read
 *                           indices of a query's hits are read with coalesced loads -- no walk at all;
 *                           per-read sample-set bitmaps (<= 128 samples): one 16-byte load per hit, OR-reduced
 *                           across the warp
 *     or, without the table (shards too large for 4 B per row):
 *     k_pop_expand + K2   per shard: rows -> LF walk (plain or to the next sampled row) -> scatter to the query's slot
 *
 * Hits are reported as  shard_id << 40 | read index  (read index = position in the shard's builder input order,
 * what the .sai gives for the walk's terminal row).
 */
#include "rs_internal.h"

#include <algorithm>

#include "rs_ops.h"
#include "rs_pop.h"

#define RS_FULL 0xFFFFFFFFu

/* ------------------------------------------------------------------------------------------------
 * read-index table: rid[row] = read index of the read whose suffix sits at `row` (0xFFFFFFFF: the
 * walk from that row does not end within max_len steps -- only possible inside the reference's tail
 * region -- or the row is not on any read's walk)
 * ---------------------------------------------------------------------------------------------- */
#define RS_RID_NONE 0xFFFFFFFFu
#define RS_RID_FAIL 0xFFFFFFFEu /* pass 3 walked from the row: the walk ends nowhere (the hit is reported as 2^64-1) */

/* pass 1: one walker per read from the row of its terminator (row i belongs to read i in builder input order);
 * every row it passes is stamped with i; term[i] = where the walk ended (lexicographic rank of the read) */
__global__ void __launch_bounds__(RS_THREADS, 4)
    k_rid_walk(const __grid_constant__ rs_index_dev ix, uint32_t max_len, uint32_t* __restrict__ rid,
               uint64_t* __restrict__ term) {
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
      if (idx < ix.n) rid[idx] = (uint32_t)i;
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      idx = nxt;
      if (sym == 0) {
        term[i] = idx;
        active = false;
      } else if (++s > max_len) {
        term[i] = ~0ull;
        active = false;
      }
    }
  }
}

/* pass 2: stamp -> read index through the walk's terminal row and the .sai */
__global__ void k_rid_resolve(uint32_t* __restrict__ rid, uint64_t n, const uint64_t* __restrict__ term,
                              const uint64_t* __restrict__ sai, uint64_t num_strings) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride) {
    const uint32_t v = rid[r];
    if (v == RS_RID_NONE) continue;
    const uint64_t t = term[v];
    uint32_t out = RS_RID_NONE;
    if (t < num_strings) {
      const uint64_t ri = sai[t];
      if (ri < RS_RID_FAIL) out = (uint32_t)ri;
    }
    rid[r] = out;
  }
}

/* pass 3: the rows passes 1-2 left open (inside the reference's tail region walks join, get cut or wander, so a row
 * there may be on no read's walk) are walked once here, exactly as the query-time fallback walks them
 * (rs_pop_walk_one), and their answer is stored: a batch then never pays a 100-step walk by one lane while the
 * other lanes of its warp wait (configs[2]: three of the 64 shards have such rows; the ranks holding them took 9-12 ms
 * for the locate stage, the others 7.3).  A read index that does not fit the table stays RS_RID_NONE (walked per query). */
__global__ void __launch_bounds__(256)
    k_rid_fill(const __grid_constant__ rs_index_dev ix, const uint64_t* __restrict__ sai, uint32_t max_steps,
               uint32_t* __restrict__ rid, unsigned long long* __restrict__ nfilled) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint32_t mine = 0;
  for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < ix.n; r += stride) {
    if (rid[r] != RS_RID_NONE) continue;
    uint64_t idx = r, res = ~0ull;
    for (uint32_t st = 0; st <= max_steps; ++st) {
      uint64_t nxt;
      const uint32_t sym = rs_lf_step(ix, idx, &nxt);
      idx = nxt;
      if (sym == 0) {
        res = idx < ix.num_strings ? sai[idx] : ~0ull;
        break;
      }
    }
    if (res == ~0ull) rid[r] = RS_RID_FAIL;
    else if (res < RS_RID_FAIL) rid[r] = (uint32_t)res;
    ++mine;
  }
  if (nfilled && mine) atomicAdd(nfilled, (unsigned long long)mine);
}

/* per-read sample-set bitmaps (<= 128 samples): bit s of read r's 16 bytes = sample s carries the read */
__global__ void k_read_sets(const uint64_t* __restrict__ samp_off, const uint8_t* __restrict__ samp_codes, uint32_t sos,
                            uint64_t num_strings, uint4* __restrict__ sets) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < num_strings; r += stride) {
    uint32_t w[4] = {0, 0, 0, 0};
    for (uint64_t t = samp_off[r]; t < samp_off[r + 1]; ++t) {
      uint32_t v = 0;
      for (uint32_t i = 0; i < sos; ++i) v = v * 223u + (uint32_t)(samp_codes[t * sos + i] - 33u);
      const uint32_t s = v - 1u; /* codes are 1-based (create_samples_hash.cpp:41-50) */
      if (s < 128u) w[s >> 5] |= 1u << (s & 31);
    }
    sets[r] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

/* ------------------------------------------------------------------------------------------------
 * pipeline kernels
 * ---------------------------------------------------------------------------------------------- */
/* packed k-mers -> their reverse complements (the RevComp task of every request, service.cpp:289-291); flagged keys
 * (bit 63: a byte outside ACGT) stay flagged */
__global__ void k_pop_revcomp(const uint64_t* __restrict__ in, uint32_t k, uint64_t Q, uint64_t* __restrict__ out) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  const uint64_t v = in[q];
  uint64_t r = 0, x = ~v;
  for (uint32_t i = 0; i < k; ++i) {
    r = (r << 2) | (x & 3u);
    x >>= 2;
  }
  out[q] = r | (v & (1ull << 63));
}

/* per query of the chunk: population count and number of hits to report */
__global__ void k_pop_totals(const uint32_t* __restrict__ cnt, uint32_t S, uint64_t C, uint64_t nq, uint64_t max_hits,
                             uint64_t* __restrict__ counts, uint64_t* __restrict__ nhits, uint32_t* __restrict__ seg,
                             uint32_t* __restrict__ pass_off, uint32_t pass_shards, const uint32_t* __restrict__ perm) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  uint64_t tot = 0, h = 0;
  for (uint32_t s = 0; s < S; ++s) {
    const uint64_t c = cnt[(uint64_t)s * C + q];
    if (seg) seg[(uint64_t)s * C + q] = (uint32_t)h; /* where shard s's hits start inside the query's slot */
    /* ... and where the shards of a gather pass start (k_pop_gather runs pass by pass) */
    if (pass_off && s % pass_shards == 0) pass_off[(uint64_t)(s / pass_shards) * C + q] = h > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)h;
    tot += c;
    h += c < max_hits ? c : max_hits;
  }
  const uint64_t qo = perm ? perm[q] : q; /* sorted batch: the answers go to the query's place in the caller's order */
  counts[qo] = tot;
  nhits[qo] = h;
}

/* hit_offsets[q] = base + local offset; the running base advances by the chunk's hits */
__global__ void k_pop_add_base(const uint64_t* __restrict__ local, uint64_t nq, const uint64_t* __restrict__ base,
                               uint64_t* __restrict__ hit_offsets) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q <= nq) hit_offsets[q] = *base + local[q]; /* entry nq = the next chunk's first (rewritten by it, same value) */
}
__global__ void k_pop_advance(const uint64_t* __restrict__ local, uint64_t nq, uint64_t* __restrict__ base,
                              uint64_t* __restrict__ hit_offsets_end, uint64_t* __restrict__ total) {
  *base += local[nq];
  if (hit_offsets_end) *hit_offsets_end = *base;
  if (total) *total = *base;
}

/* extractPrefix from one row of shard `v` (rows the read-index table does not cover): as K2 walks it.  A copy per
 * instantiation of the gather kernel: ptxas 12.9 crashes on one out-of-line function called from several of them. */
template <uint32_t G>
__device__ __noinline__ uint64_t rs_pop_walk_one(const rs_index_dev* __restrict__ views, uint32_t s,
                                                 const uint64_t* const* __restrict__ sais, uint64_t row,
                                                 uint32_t max_steps) {
  const rs_index_dev& ix = views[s]; /* fields read from global memory as needed: this path is rare */
  uint64_t idx = row;
  for (uint32_t st = 0; st <= max_steps; ++st) {
    uint64_t nxt;
    const uint32_t sym = rs_lf_step(ix, idx, &nxt);
    idx = nxt;
    if (sym == 0) return idx < ix.num_strings ? sais[s][idx] : ~0ull;
  }
  return ~0ull;
}

/* rank owning the reduced result of a packed k-mer: its first four bases as a base-4 number modulo the ranks
 * (sharding.owner_of; same rule as K1's fused count reduction, rs_kernels.cu) */
__device__ __forceinline__ uint32_t rs_owner_of(uint64_t pk, uint32_t k, bool, uint32_t world) {
  const uint32_t prefix = k >= 4 ? (uint32_t)(pk >> (2 * (k - 4))) & 0xFFu : (uint32_t)pk & 0xFFu;
  return prefix % world;
}

/* multi-GPU: per query of the chunk, this rank's count is added into the owner's population count and its number of
 * hits stored into the owner's per-rank table -- both straight into the owner's memory over NVLink */
__global__ void k_pop_totals_comm(const __grid_constant__ rs_pop_peers P, const uint64_t* __restrict__ keys, uint32_t k,
                                  const uint32_t* __restrict__ cnt, uint32_t S, uint64_t C, uint64_t q0, uint64_t nq,
                                  uint64_t max_hits, int want_hits, uint32_t* __restrict__ pass_off, uint32_t pass_shards,
                                  const uint32_t* __restrict__ perm) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  uint64_t tot = 0, h = 0;
  for (uint32_t s = 0; s < S; ++s) {
    const uint64_t c = cnt[(uint64_t)s * C + q];
    if (pass_off && s % pass_shards == 0) pass_off[(uint64_t)(s / pass_shards) * C + q] = h > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)h;
    tot += c;
    h += c < max_hits ? c : max_hits;
  }
  const uint32_t owner = rs_owner_of(keys[q], k, false, P.world);
  const uint64_t qo = q0 + (perm ? perm[q] : q);
  if (tot) atomicAdd(reinterpret_cast<unsigned long long*>(P.counts[owner]) + qo, (unsigned long long)tot);
  if (want_hits) P.rankh[owner][(uint64_t)P.rank * P.maxq + qo] = h > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)h;
}

/* owner: hits of each query it owns, summed over the source ranks (0 for the queries of other owners) */
__global__ void k_pop_owner_hits(const __grid_constant__ rs_pop_peers P, const uint64_t* __restrict__ keys, uint32_t k,
                                 uint64_t q0, uint64_t nq, uint64_t* __restrict__ nhits) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  uint64_t h = 0;
  if (rs_owner_of(keys[q], k, false, P.world) == P.rank)
    for (uint32_t r = 0; r < P.world; ++r) h += P.rankh[P.rank][(uint64_t)r * P.maxq + q0 + q];
  nhits[q] = h;
}

/* owner: tells every source rank where its hits of each owned query go (rank 0's first, then rank 1's, ...) */
__global__ void k_pop_owner_slots(const __grid_constant__ rs_pop_peers P, const uint64_t* __restrict__ keys, uint32_t k,
                                  uint64_t q0, uint64_t nq, const uint64_t* __restrict__ hit_offsets) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  if (rs_owner_of(keys[q], k, false, P.world) != P.rank) return;
  uint64_t run = hit_offsets[q0 + q];
  for (uint32_t r = 0; r < P.world; ++r) {
    P.slot[r][q0 + q] = run;
    run += P.rankh[P.rank][(uint64_t)r * P.maxq + q0 + q];
  }
}

/* One warp per query: its hits shard after shard, rows in order, straight from the read-index tables; the union of
 * the hit reads' sample sets when carrier sets are wanted (W = 4 words of 32 samples).  The hits go to the rank that
 * owns the query (P.world > 1: over NVLink peer memory, at the slot the owner assigned to this rank -- locate and the
 * fan-in of the hit lists, server.cpp:184-197, are one kernel), the sample set is OR-ed into the owner's bitmap. */
#ifndef RS_POP_SHARD_MAJOR_MIN
#define RS_POP_SHARD_MAJOR_MIN (1u << 17) /* queries per chunk from which the gather runs shard by shard */
#endif
#ifndef RS_POP_PASS_BYTES
#define RS_POP_PASS_BYTES (32ull << 20)
#endif
#define RS_GATHER_LANE_ROWS 8u /* intervals up to this many rows are read by one lane; longer ones by the whole warp */
#ifndef RS_GATHER_BLOCKS
#define RS_GATHER_BLOCKS 4 /* resident CTAs per SM: the kernel waits on memory, occupancy is its parallelism */
#endif
/* OR over the G lanes of a query's group (every lane of the warp takes part) */
template <uint32_t G>
__device__ __forceinline__ uint32_t rs_group_or(uint32_t v) {
  if constexpr (G == 32) {
    return __reduce_or_sync(RS_FULL, v);
  } else {
#pragma unroll
    for (int d = (int)G / 2; d >= 1; d >>= 1) v |= __shfl_xor_sync(RS_FULL, v, d, G);
    return v;
  }
}
/* G lanes per query (a power of two, 1..32), 32 / G queries per warp, the lanes of a group taking one shard each, G
 * shards per round.  The kernel waits on memory and its parallelism is the lanes that have a load in flight: a rank that
 * holds few shards (8 or 9 of the 64 on each of 8 GPUs) would leave most lanes of a warp-per-query kernel without a
 * shard, so the host picks the largest G that keeps >= 85 % of the lane slots on a shard (rs_gather_lanes).  All groups
 * of a warp run the same trips (a group past the batch's end carries zero counts), so the warp stays converged and an
 * interval of many rows is read by all 32 lanes whichever group it belongs to. */
template <uint32_t G>
__global__ void __launch_bounds__(256, RS_GATHER_BLOCKS)
    k_pop_gather(const __grid_constant__ rs_pop_tables T, const __grid_constant__ rs_pop_peers P,
                 const uint64_t* __restrict__ keys, uint32_t k, const uint64_t* __restrict__ lo,
                 const uint32_t* __restrict__ cnt, uint64_t C, uint64_t q0, uint64_t nq, uint64_t max_hits,
                 const uint64_t* __restrict__ slot, uint64_t capacity, int want_carriers, uint32_t s_begin,
                 uint32_t s_end, const uint32_t* __restrict__ pass_off) {
  constexpr uint32_t QW = 32u / G; /* queries per warp */
  const uint32_t wlane = threadIdx.x & 31;
  const uint32_t lane = wlane & (G - 1u); /* lane within the query's group */
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t qb = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * QW; qb < nq; qb += warps * QW) {
    const uint64_t q = qb + wlane / G;
    const bool valid = q < nq;
    const uint32_t owner = (valid && P.world > 1) ? rs_owner_of(keys[q], k, false, P.world) : 0u;
    uint64_t* __restrict__ hits = P.hits[owner];
    uint64_t off = valid ? slot[q0 + q] + (pass_off ? (uint64_t)pass_off[q] : 0ull) : 0;
    uint32_t b0 = 0, b1 = 0, b2 = 0, b3 = 0;
    /* The table reads below are 4 and 16 bytes at random addresses, and HBM delivers whole 128-byte lines for them
     * (ncu on configs[2]: 247 GB of DRAM reads per 10 M-query step for 790 M hits = 1.9 G line fetches at 26 G lines/s,
     * the rate of the random-access probe); ld.global.cg instead of the L1 path made it slower (82 vs 73 ms).
     * A k-mer of a read set at 30x hits a handful of rows in most of the shards: the lanes take one shard each
     * (32 independent chains of loads in flight per warp), a scan over the group lays their rows out in shard order. */
    for (uint32_t s0 = s_begin; s0 < s_end; s0 += G) {
      const uint32_t s = s0 + lane;
      uint64_t c = (valid && s < s_end) ? cnt[(uint64_t)s * C + q] : 0;
      if (c > max_hits) c = max_hits;
      uint64_t incl = c;
#pragma unroll
      for (int d = 1; d < (int)G; d <<= 1) {
        const uint64_t t = __shfl_up_sync(RS_FULL, incl, d, G);
        if ((int)lane >= d) incl += t;
      }
      const uint64_t total = __shfl_sync(RS_FULL, incl, G - 1, G);
      const uint64_t mine = off + incl - c; /* where this lane's shard starts in the query's list */
      const uint64_t first = c ? lo[(uint64_t)s * C + q] : 0;
      if (c && c <= RS_GATHER_LANE_ROWS) {
        const uint32_t* __restrict__ rid = T.rid[s];
        const uint4* __restrict__ sets = T.sets[s];
        const uint64_t tag = T.tag[s];
        for (uint32_t j = 0; j < (uint32_t)c; ++j) {
          uint64_t r = rid[first + j];
          if (r == RS_RID_NONE) r = rs_pop_walk_one<G>(T.views, s, T.sai, first + j, T.max_steps);
          else if (r == RS_RID_FAIL) r = ~0ull;
          if (mine + j < capacity) hits[mine + j] = r == ~0ull ? ~0ull : (tag | r);
          if (want_carriers && sets && r != ~0ull) {
            const uint4 v = __ldg(sets + r);
            b0 |= v.x;
            b1 |= v.y;
            b2 |= v.z;
            b3 |= v.w;
          }
        }
      }
      /* long intervals (a repeat, a short k): the whole warp reads the rows of one shard, coalesced */
      uint32_t big = __ballot_sync(RS_FULL, c > RS_GATHER_LANE_ROWS);
      while (big) {
        const int src = __ffs((int)big) - 1;
        big &= big - 1;
        const uint64_t bc = __shfl_sync(RS_FULL, c, src), bfirst = __shfl_sync(RS_FULL, first, src),
                       bmine = __shfl_sync(RS_FULL, mine, src);
        const uint32_t bs = __shfl_sync(RS_FULL, s, src);
        uint64_t* __restrict__ bhits = P.hits[__shfl_sync(RS_FULL, owner, src)];
        const uint32_t* __restrict__ rid = T.rid[bs];
        const uint4* __restrict__ sets = T.sets[bs];
        const uint64_t tag = T.tag[bs];
        uint32_t t0 = 0, t1 = 0, t2 = 0, t3 = 0;
        for (uint64_t j = wlane; j < bc; j += 32) {
          uint64_t r = rid[bfirst + j];
          if (r == RS_RID_NONE) r = rs_pop_walk_one<G>(T.views, bs, T.sai, bfirst + j, T.max_steps);
          else if (r == RS_RID_FAIL) r = ~0ull;
          if (bmine + j < capacity) bhits[bmine + j] = r == ~0ull ? ~0ull : (tag | r);
          if (want_carriers && sets && r != ~0ull) {
            const uint4 v = __ldg(sets + r);
            t0 |= v.x;
            t1 |= v.y;
            t2 |= v.z;
            t3 |= v.w;
          }
        }
        if (want_carriers) { /* the set goes to the query of the lane the interval came from */
          t0 = __reduce_or_sync(RS_FULL, t0);
          t1 = __reduce_or_sync(RS_FULL, t1);
          t2 = __reduce_or_sync(RS_FULL, t2);
          t3 = __reduce_or_sync(RS_FULL, t3);
          if ((int)wlane == src) {
            b0 |= t0;
            b1 |= t1;
            b2 |= t2;
            b3 |= t3;
          }
        }
      }
      off += total;
    }
    if (want_carriers) {
      b0 = rs_group_or<G>(b0);
      b1 = rs_group_or<G>(b1);
      b2 = rs_group_or<G>(b2);
      b3 = rs_group_or<G>(b3);
      if (valid) {
        uint32_t* dst = P.carr[owner] + (q0 + q) * 4;
        if (P.world == 1 && s_begin == 0) { /* the first pass stores, the others add */
          if (lane == 0) *reinterpret_cast<uint4*>(dst) = make_uint4(b0, b1, b2, b3);
        } else { /* several ranks contribute to one bitmap (cleared by its owner before the step) */
          for (uint32_t w = lane; w < 4; w += G) {
            const uint32_t v = w == 0 ? b0 : w == 1 ? b1 : w == 2 ? b2 : b3;
            if (v) atomicOr(dst + w, v);
          }
        }
      }
    }
  }
}

/* cold path of the shard-major gather: the rows of this lane's interval that the table leaves to a walk */
__device__ __noinline__ void rs_gather_open_rows(const rs_pop_tables& T, uint32_t s, const uint32_t* __restrict__ rid,
                                                 const uint4* __restrict__ sets, uint64_t tag, uint32_t c, uint64_t first,
                                                 uint64_t* __restrict__ dst, uint32_t* bm) {
  for (uint32_t j = 0; j < c; ++j) {
    if (rid[first + j] != RS_RID_NONE) continue;
    const uint64_t r = rs_pop_walk_one<0>(T.views, s, T.sai, first + j, T.max_steps);
    if (r == ~0ull) continue; /* the hot loop stored 2^64-1 */
    dst[j] = tag | r;
    if (sets) {
      const uint4 v = __ldg(sets + r);
      bm[0] |= v.x;
      bm[1] |= v.y;
      bm[2] |= v.z;
      bm[3] |= v.w;
    }
  }
}

/* Shard-major gather: ONE shard per launch.  A warp takes 32 consecutive queries of the (sorted) chunk per trip: their
 * counts, first rows and destinations are read coalesced (28 bytes per query, streamed), then the rows of all 32
 * intervals together are dealt out to the lanes -- hit h of the trip to lane h mod 32, its query found by a search
 * over the warp's running sum in shared memory -- so a trip costs ceil(hits / 32) rounds of
 * row -> read index -> sample set -> store, whatever the split of the hits over the queries (one lane per query ran
 * as many rounds as the longest interval of the 32, with the other lanes masked: ncu showed 13 of 32 lanes active and
 * 414 warp instructions per trip).  An interval of a million rows is simply 31 k rounds of coalesced reads.  A query's
 * hits of this shard go to slot[q] + seg[q] (seg: the hits of the shards before this one, from k_pop_totals); sample
 * sets are OR-ed per query in shared memory, then into the chunk's bitmap by the lane that owns the query (launches
 * follow each other on the stream: no global atomics).  With a sorted batch the rows of successive queries ascend, so
 * the read-index table streams through L2 once per launch. */
#ifndef RS_GATHER_SHARD_BLOCKS
#define RS_GATHER_SHARD_BLOCKS 4
#endif
__global__ void __launch_bounds__(256, RS_GATHER_SHARD_BLOCKS)
    k_pop_gather_shard(const __grid_constant__ rs_pop_tables T, const __grid_constant__ rs_pop_peers P, uint32_t s,
                       const uint64_t* __restrict__ keys, uint32_t k, const uint64_t* __restrict__ lo_s,
                       const uint32_t* __restrict__ cnt_s, const uint32_t* __restrict__ seg_s, uint64_t q0, uint64_t nq,
                       uint64_t max_hits, const uint64_t* __restrict__ slot, uint64_t capacity, uint4* __restrict__ carr) {
  __shared__ uint64_t sh_incl[8][32];  /* running sum of the trip's counts */
  __shared__ uint64_t sh_first[8][32]; /* first row of each query's interval */
  __shared__ uint64_t sh_dst[8][32];   /* address of each query's first hit in its owner's list */
  __shared__ uint32_t sh_bm[8][32][4]; /* sample sets of the trip's queries */
  const uint32_t wlane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
  const uint32_t* __restrict__ rid = T.rid[s];
  const uint4* __restrict__ sets = carr ? T.sets[s] : nullptr;
  const uint64_t tag = T.tag[s];
  const bool multi = P.world > 1;
  for (uint64_t qb = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; qb < nq; qb += nthreads) { /* warp-uniform */
    const uint64_t q = qb + wlane;
    uint64_t c = q < nq ? cnt_s[q] : 0;
    if (c > max_hits) c = max_hits;
    if (!__any_sync(RS_FULL, c != 0)) continue;
    const uint64_t first = c ? lo_s[q] : 0;
    const uint64_t mine = c ? slot[q0 + q] + seg_s[q] : 0;
    const uint32_t owner = (c && multi) ? rs_owner_of(keys[q], k, false, P.world) : 0u;
    if (mine >= capacity) c = 0;                       /* nothing of it can be reported */
    else if (c > capacity - mine) c = capacity - mine; /* the list ends inside the interval */
    uint64_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t t = __shfl_up_sync(RS_FULL, incl, d);
      if ((int)wlane >= d) incl += t;
    }
    const uint64_t total = __shfl_sync(RS_FULL, incl, 31);
    sh_incl[w][wlane] = incl;
    sh_first[w][wlane] = first;
    sh_dst[w][wlane] = (uint64_t)(uintptr_t)(P.hits[owner] + mine);
    if (sets) *reinterpret_cast<uint4*>(sh_bm[w][wlane]) = make_uint4(0, 0, 0, 0);
    __syncwarp();
    bool open_rows = false;
    for (uint64_t hb = 0; hb < total; hb += 32) {
      const uint64_t h = hb + wlane;
      const bool live = h < total;
      uint32_t l = 0; /* the query hit h belongs to: the first whose running sum exceeds h */
      uint4 sv = make_uint4(0, 0, 0, 0);
      if (live) {
#pragma unroll
        for (uint32_t step = 16; step; step >>= 1)
          if (sh_incl[w][l + step - 1] <= h) l += step;
        const uint64_t j = h - (l ? sh_incl[w][l - 1] : 0ull);
        const uint32_t v = rid[sh_first[w][l] + j];
        open_rows |= v == RS_RID_NONE; /* left to a walk: redone below */
        reinterpret_cast<uint64_t*>((uintptr_t)sh_dst[w][l])[j] = v >= RS_RID_FAIL ? ~0ull : (tag | v);
        if (sets && v < RS_RID_FAIL) sv = __ldg(sets + v);
      }
      if (sets) {
        __syncwarp();
        const uint32_t l0 = __shfl_sync(RS_FULL, l, 0);
        if (__all_sync(RS_FULL, !live || l == l0)) { /* one interval fills the round: one OR for the warp */
          sv.x = __reduce_or_sync(RS_FULL, sv.x);
          sv.y = __reduce_or_sync(RS_FULL, sv.y);
          sv.z = __reduce_or_sync(RS_FULL, sv.z);
          sv.w = __reduce_or_sync(RS_FULL, sv.w);
          if (wlane == 0) {
            sh_bm[w][l0][0] |= sv.x;
            sh_bm[w][l0][1] |= sv.y;
            sh_bm[w][l0][2] |= sv.z;
            sh_bm[w][l0][3] |= sv.w;
          }
          __syncwarp();
        } else {
          if (sv.x) atomicOr(&sh_bm[w][l][0], sv.x);
          if (sv.y) atomicOr(&sh_bm[w][l][1], sv.y);
          if (sv.z) atomicOr(&sh_bm[w][l][2], sv.z);
          if (sv.w) atomicOr(&sh_bm[w][l][3], sv.w);
        }
      }
    }
    __syncwarp();
    if (__any_sync(RS_FULL, open_rows)) /* rows whose read index does not fit the table: walked, one lane per query */
      rs_gather_open_rows(T, s, rid, sets, tag, (uint32_t)c, first, reinterpret_cast<uint64_t*>((uintptr_t)sh_dst[w][wlane]),
                          sh_bm[w][wlane]);
    if (sets) {
      __syncwarp();
      const uint4 b = *reinterpret_cast<const uint4*>(sh_bm[w][wlane]);
      if (b.x | b.y | b.z | b.w) {
        uint4 v = carr[q];
        v.x |= b.x;
        v.y |= b.y;
        v.z |= b.z;
        v.w |= b.w;
        carr[q] = v;
      }
    }
    __syncwarp(); /* the next trip rewrites the shared arrays */
  }
}

/* several GPUs: the chunk's locally united sample sets are OR-ed into the owners' bitmaps (cleared by them before the step) */
__global__ void k_pop_carr_send(const __grid_constant__ rs_pop_peers P, const uint64_t* __restrict__ keys, uint32_t k,
                                uint64_t q0, uint64_t nq, const uint4* __restrict__ carr, const uint32_t* __restrict__ perm) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  const uint4 v = carr[q];
  if (!(v.x | v.y | v.z | v.w)) return;
  uint32_t* dst = P.carr[rs_owner_of(keys[q], k, false, P.world)] + (q0 + (perm ? perm[q] : q)) * 4;
  if (v.x) atomicOr(dst + 0, v.x);
  if (v.y) atomicOr(dst + 1, v.y);
  if (v.z) atomicOr(dst + 2, v.z);
  if (v.w) atomicOr(dst + 3, v.w);
}

/* sorted batch: 0, 1, 2, ... (the values the sort carries along); where each sorted query's hits start; the sample
 * sets back in the caller's order */
__global__ void k_pop_iota(uint32_t* __restrict__ out, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)i;
}
/* sort keys of a batch whose answers are to become final range by range of the caller's order: the range (i / per) above
 * the 2k bits of the k-mer; and the batch's keys in the sorted order (the sort keys drop the flag bit) */
__global__ void k_pop_range_keys(const uint64_t* __restrict__ keys, uint64_t n, uint64_t per, uint32_t bits,
                                 uint64_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = ((i / per) << bits) | (keys[i] & ((1ull << bits) - 1));
}
__global__ void k_pop_keys_sorted(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ perm, uint64_t n,
                                  uint64_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[perm[i]];
}
__global__ void k_pop_slot_sorted(const uint64_t* __restrict__ slot, const uint32_t* __restrict__ perm, uint64_t n,
                                  uint64_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = slot[perm[i]];
}
__global__ void k_pop_carr_unsort(const uint4* __restrict__ carr, const uint32_t* __restrict__ perm, uint64_t n,
                                  uint4* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[perm[i]] = carr[i];
}

/* hits [begin, end) of the list as 40-bit values (rs_pop_query_hits40): shard id << 32 | read index, little-endian,
 * 5 bytes each; 2^40 - 1 = no read.  Eight hits = five 64-bit words: a thread packs one aligned group of eight, the
 * groups the range cuts are written byte by byte.  *bad is set when a shard id >= 255 or a read index >= 2^32 appears. */
__global__ void k_pop_pack40(const uint64_t* __restrict__ hits, const uint64_t* __restrict__ begin_dev,
                             const uint64_t* __restrict__ end_dev, uint64_t capacity, uint8_t* __restrict__ out,
                             uint32_t* __restrict__ bad) {
  const uint64_t b = begin_dev ? (*begin_dev < capacity ? *begin_dev : capacity) : 0;
  const uint64_t e = *end_dev < capacity ? *end_dev : capacity;
  const uint64_t g0 = b / 8, g1 = (e + 7) / 8;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  bool any_bad = false;
  for (uint64_t g = g0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < g1; g += stride) {
    const uint64_t h0 = g * 8;
    uint64_t v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint64_t at = h0 + i;
      uint64_t h = (at >= b && at < e) ? hits[at] : 0;
      if (h == ~0ull) h = (1ull << 40) - 1;
      else {
        any_bad |= (h >> 40) >= 255u || ((h >> 32) & 0xFFu) != 0;
        h = ((h >> 40) << 32) | (h & 0xFFFFFFFFull);
      }
      v[i] = h;
    }
    if (h0 >= b && h0 + 8 <= e) {
      uint64_t* o = reinterpret_cast<uint64_t*>(out + h0 * 5); /* 40 bytes, 8-byte aligned */
      o[0] = v[0] | (v[1] << 40);
      o[1] = (v[1] >> 24) | (v[2] << 16) | (v[3] << 56);
      o[2] = (v[3] >> 8) | (v[4] << 32);
      o[3] = (v[4] >> 32) | (v[5] << 8) | (v[6] << 48);
      o[4] = (v[6] >> 16) | (v[7] << 24);
    } else {
      for (int i = 0; i < 8; ++i) {
        const uint64_t at = h0 + i;
        if (at < b || at >= e) continue;
        for (int x = 0; x < 5; ++x) out[at * 5 + x] = (uint8_t)(v[i] >> (8 * x));
      }
    }
  }
  if (any_bad) *bad = 1u;
}

/* lanes per query for a rank holding S shards: the largest power of two that keeps >= 85 % of the lane slots of a round
 * on a shard (S = 64, 32: 32; 16: 16; 8, 7: 8; 21, 22: 8; 9: 2; 1 lane always qualifies) */
static uint32_t rs_gather_lanes(uint32_t S) {
  for (uint32_t g = 32; g > 1; g >>= 1) {
    const uint32_t slots = (S + g - 1) / g * g;
    if ((uint64_t)S * 100 >= (uint64_t)slots * 85) return g;
  }
  return 1;
}

/* ---- without the read-index table: per shard, rows -> walk -> scatter ---- */
__global__ void k_pop_shard_counts(const uint32_t* __restrict__ cnt_s, uint64_t nq, uint64_t max_hits,
                                   uint64_t* __restrict__ out) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  const uint64_t c = cnt_s[q];
  out[q] = c < max_hits ? c : max_hits;
}

/* rows[loff[q] + j] = first[q] + j,  dst[..] = ADDRESS of the hit's slot in the owner's list (0: beyond its capacity);
 * one warp per query */
__global__ void k_pop_expand(const __grid_constant__ rs_pop_peers P, const uint64_t* __restrict__ keys, uint32_t k,
                             const uint64_t* __restrict__ first, const uint64_t* __restrict__ loff, uint64_t q0, uint64_t nq,
                             const uint64_t* __restrict__ slot, const uint32_t* __restrict__ seg, uint64_t bound,
                             uint64_t capacity, uint64_t* __restrict__ rows, uint64_t* __restrict__ dst) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < nq; q += warps) {
    const uint64_t b = loff[q], e = loff[q + 1];
    if (b == e) continue;
    const uint32_t owner = P.world > 1 ? rs_owner_of(keys[q], k, false, P.world) : 0u;
    const uint64_t f = first[q], d0 = slot[q0 + q] + seg[q];
    for (uint64_t j = b + lane; j < e && j < bound; j += 32) { /* rows past `bound` could not be reported anyway */
      rows[j] = f + (j - b);
      const uint64_t d = d0 + (j - b);
      dst[j] = d < capacity ? (uint64_t)(uintptr_t)(P.hits[owner] + d) : 0ull;
    }
  }
}

/* *dst[i] = tag | read_index[i]; n read from the device (the shard's hit count of this chunk) */
__global__ void k_pop_scatter(const uint64_t* __restrict__ read_index, const uint64_t* __restrict__ dst,
                              const uint64_t* __restrict__ n_dev, uint64_t bound, uint64_t tag) {
  const uint64_t n = *n_dev < bound ? *n_dev : bound;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t r = read_index[i];
    uint64_t* d = reinterpret_cast<uint64_t*>((uintptr_t)dst[i]);
    if (d) *d = r == ~0ull ? ~0ull : (tag | r);
  }
}

/* union of the hit reads' sample sets from the hit list (any number of samples: a bitmap per warp in shared memory;
 * sample codes decoded as create_samples_hash.cpp:31-51 wrote them) */
__global__ void __launch_bounds__(256)
    k_pop_carriers(const __grid_constant__ rs_pop_tables T, const uint64_t* __restrict__ hit_offsets, uint64_t nq,
                   const uint64_t* __restrict__ hits, uint64_t capacity, uint32_t words, uint32_t* __restrict__ carriers) {
  /* single-GPU sets only: the sample tables of a hit's shard must be resident on the rank that holds the hit list */
  extern __shared__ uint32_t smem[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  uint32_t* bm = smem + wib * words;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < nq; q += warps) {
    for (uint32_t w = lane; w < words; w += 32) bm[w] = 0;
    __syncwarp();
    const uint64_t b = hit_offsets[q], e = hit_offsets[q + 1] < capacity ? hit_offsets[q + 1] : capacity;
    for (uint64_t j = b + lane; j < e; j += 32) {
      const uint64_t h = hits[j];
      if (h == ~0ull) continue;
      const uint32_t s = T.slot_of_shard[h >> 40];
      const uint64_t r = h & ((1ull << 40) - 1);
      const uint64_t* __restrict__ so = T.samp_off[s];
      const uint8_t* __restrict__ sc = T.samp_codes[s];
      if (!so) continue;
      for (uint64_t t = so[r]; t < so[r + 1]; ++t) {
        uint32_t v = 0;
        for (uint32_t i = 0; i < T.sos; ++i) v = v * 223u + (uint32_t)(sc[t * T.sos + i] - 33u);
        const uint32_t smp = v - 1u;
        if (smp < words * 32u) atomicOr(&bm[smp >> 5], 1u << (smp & 31));
      }
    }
    __syncwarp();
    for (uint32_t w = lane; w < words; w += 32) carriers[q * words + w] = bm[w];
    __syncwarp();
  }
}

/* ------------------------------------------------------------------------------------------------
 * host side
 * ---------------------------------------------------------------------------------------------- */
struct rs_shardset {
  int device = 0;
  bool sealed = false;
  std::vector<rs_index*> shards;
  std::vector<uint32_t> shard_ids;
  std::vector<uint32_t*> d_rid;   /* owned: read-index tables */
  std::vector<uint4*> d_sets;     /* owned: per-read sample-set bitmaps */
  uint32_t max_read_len = 0, locate_every = 0;
  uint32_t num_samples = 0, words = 0; /* carrier-set width */
  int sos = 0;
  bool have_samples = false, have_sets = false;
  uint64_t table_bytes = 0;
  rs_pop_tables tab;              /* device pointer tables (point into d_tabmem) */
  void* d_tabmem = nullptr;
  /* per-call scratch, grown on demand; calls on one set are serialised */
  std::mutex mu;
  cudaStream_t stream = nullptr;
  Scratch s_ckeys, s_sort, s_skeys, s_perm, s_iota, s_sslot, s_carrl, s_pass, s_packed, s_lo, s_cnt, s_seg, s_nh, s_loc, s_scan, s_base, s_rows, s_dst, s_ri, s_in, s_lcnt;
  /* host-pointer calls: device-side result buffers, the copy-out stream, per-sub-batch events and hit-list marks */
  Scratch s_hc, s_ho, s_hh, s_hw, s_hm, s_h40;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_sub[RS_POP_MAX_SUB] = {};
  uint64_t* h_marks = nullptr;
  uint64_t* d_stats = nullptr; /* executed updateInterval steps / LF steps of the set's calls (bench.py) */
  uint64_t chunk = 1u << 20;
  uint32_t gather_lanes = 32; /* lanes per query in the gather kernel (rs_gather_lanes; RS_B200_GATHER_LANES forces) */
  bool sort_batches = true;   /* RS_B200_POP_SORT=0 turns the sorted batch off */
  int gather_mode = 0;        /* 0: by batch size; 1: query-major passes (k_pop_gather); 2: shard-major (k_pop_gather_shard) */
  uint32_t pass_shards = 0;   /* shards per gather pass (0 = all in one): a pass's sample-set tables fit L2 */
  bool chunk_forced = false; /* RS_B200_POP_CHUNK given */
  /* stage timing (bench.py: which kernel a step's time goes to): events recorded on the call's stream around the
   * stages of every chunk, read back -- and summed per stage -- by rs_shardset_read_timing */
  bool timing = false;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_used = 0;
  struct Span { int stage; size_t a, b; };
  std::vector<Span> spans;
  uint64_t launches[RS_POP_STAGES] = {0, 0, 0, 0};
  size_t mark(cudaStream_t st) {
    if (ev_used == ev_pool.size()) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) return (size_t)-1;
      ev_pool.push_back(e);
    }
    cudaEventRecord(ev_pool[ev_used], st);
    return ev_used++;
  }
};

/* RAII: one timed span of a stage on the stream (no-op unless timing is on) */
struct StageTimer {
  rs_shardset* set;
  cudaStream_t st;
  int stage;
  size_t a;
  StageTimer(rs_shardset* s, cudaStream_t stream, int stg, uint64_t nlaunch) : set(s), st(stream), stage(stg), a((size_t)-1) {
    if (set->timing && set->spans.size() < 100000) {
      a = set->mark(st);
      set->launches[stage] += nlaunch;
    }
  }
  ~StageTimer() {
    if (a != (size_t)-1) {
      const size_t b = set->mark(st);
      if (b != (size_t)-1) set->spans.push_back({stage, a, b});
    }
  }
};

int rs_shardset_create(int device, rs_shardset** out) {
  if (!out) return fail(RS_STATUS_ARG, "out is null");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(RS_STATUS_CUDA, "CUDA device %d not available (%d devices): this engine has no CPU path", device, ndev);
  CU(cudaSetDevice(device));
  rs_shardset* s = new rs_shardset;
  s->device = device;
  memset(&s->tab, 0, sizeof(s->tab));
  const char* e = getenv("RS_B200_POP_CHUNK");
  if (e && atoll(e) >= 1024) {
    s->chunk = (uint64_t)atoll(e);
    s->chunk_forced = true;
  }
  cudaError_t ce = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  if (ce != cudaSuccess) {
    delete s;
    return fail(RS_STATUS_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(ce));
  }
  *out = s;
  return RS_STATUS_OK;
}

int rs_shardset_add(rs_shardset* set, rs_index* ix, uint32_t shard_id) {
  if (!set || !ix) return fail(RS_STATUS_ARG, "null argument");
  if (set->sealed) return fail(RS_STATUS_ARG, "shard set is sealed");
  if (ix->device != set->device) return fail(RS_STATUS_ARG, "shard lives on another device");
  if (shard_id >= RS_POP_MAX_SHARD_ID) return fail(RS_STATUS_ARG, "shard id must be below %u", RS_POP_MAX_SHARD_ID);
  if (set->shards.size() >= RS_POP_MAX_SHARDS) return fail(RS_STATUS_TOO_LARGE, "at most %d shards per set", RS_POP_MAX_SHARDS);
  if (!ix->d_sai) return fail(RS_STATUS_NO_TABLE, "shard %u has no .sai table attached", shard_id);
  if (ix->dev.num_strings >> 32) return fail(RS_STATUS_TOO_LARGE, "shard %u holds more than 2^32 reads", shard_id);
  for (uint32_t id : set->shard_ids)
    if (id == shard_id) return fail(RS_STATUS_ARG, "shard id %u added twice", shard_id);
  set->shards.push_back(ix);
  set->shard_ids.push_back(shard_id);
  return RS_STATUS_OK;
}

static void free_tables(rs_shardset* set) {
  for (auto p : set->d_rid)
    if (p) cudaFree(p);
  for (auto p : set->d_sets)
    if (p) cudaFree(p);
  set->d_rid.clear();
  set->d_sets.clear();
  if (set->d_tabmem) cudaFree(set->d_tabmem);
  set->d_tabmem = nullptr;
  set->table_bytes = 0;
}

void rs_shardset_destroy(rs_shardset* set) {
  if (!set) return;
  cudaSetDevice(set->device);
  if (set->stream) cudaStreamSynchronize(set->stream);
  free_tables(set);
  Scratch* all[] = {&set->s_ckeys, &set->s_sort, &set->s_skeys, &set->s_perm, &set->s_iota, &set->s_sslot, &set->s_carrl, &set->s_pass, &set->s_packed, &set->s_lo, &set->s_cnt, &set->s_seg, &set->s_nh, &set->s_loc, &set->s_scan,
                    &set->s_base, &set->s_rows, &set->s_dst, &set->s_ri, &set->s_in, &set->s_lcnt,
                    &set->s_hc, &set->s_ho, &set->s_hh, &set->s_hw, &set->s_hm, &set->s_h40};
  for (Scratch* s : all) s->release();
  if (set->copy_stream) cudaStreamDestroy(set->copy_stream);
  if (set->h_marks) cudaFreeHost(set->h_marks);
  for (int i = 0; i < RS_POP_MAX_SUB; ++i)
    if (set->ev_sub[i]) cudaEventDestroy(set->ev_sub[i]);
  if (set->d_stats) cudaFree(set->d_stats);
  for (cudaEvent_t e : set->ev_pool) cudaEventDestroy(e);
  if (set->stream) cudaStreamDestroy(set->stream);
  delete set;
}

static int grid_for(uint64_t items, int per_block, int cap_blocks) {
  uint64_t b = (items + per_block - 1) / per_block;
  if (b > (uint64_t)cap_blocks) b = cap_blocks;
  return b ? (int)b : 1;
}

int rs_shardset_seal(rs_shardset* set, uint32_t max_read_len, uint32_t locate_every) {
  if (!set) return fail(RS_STATUS_ARG, "null argument");
  if (set->shards.empty()) return fail(RS_STATUS_ARG, "shard set is empty");
  if (max_read_len < 1 || max_read_len >= (1u << 24) - 1) return fail(RS_STATUS_ARG, "max_read_len must be in [1, 2^24 - 2]");
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  CU(cudaStreamSynchronize(set->stream));
  free_tables(set);
  set->sealed = false;
  const size_t S = set->shards.size();
  cudaStream_t st = set->stream;
  /* sample tables: all shards or none, one code width */
  set->have_samples = true;
  set->sos = 0;
  set->num_samples = 0;
  for (rs_index* ix : set->shards) {
    if (!ix->d_samp_off) set->have_samples = false;
  }
  if (set->have_samples) {
    set->sos = set->shards[0]->size_of_sample;
    for (rs_index* ix : set->shards) {
      if (ix->size_of_sample != set->sos) return fail(RS_STATUS_ARG, "shards disagree on size_of_sample");
      set->num_samples = std::max(set->num_samples, ix->num_samples);
    }
  }
  set->words = set->have_samples ? std::max<uint32_t>(4u, ((set->num_samples + 31) / 32 + 3) / 4 * 4) : 0;
  set->have_sets = set->have_samples && set->num_samples <= 128 && locate_every == 1;
  if (set->words > 1536) return fail(RS_STATUS_TOO_LARGE, "carrier sets support at most 49152 samples (tables hold %u)", set->num_samples);
  set->max_read_len = max_read_len;
  set->locate_every = locate_every;

  set->d_rid.assign(S, nullptr);
  set->d_sets.assign(S, nullptr);
  if (locate_every == 1) {
    /* read-index tables, one shard at a time (the walkers of a shard keep its blocks in L2) */
    for (size_t i = 0; i < S; ++i) {
      rs_index* ix = set->shards[i];
      rs_index_dev v = ix->dev;
      v.loc = v.loc_fin = v.loc_mark = v.loc_rank = nullptr;
      v.loc_every = 0;
      const uint64_t n = v.n, ns = v.num_strings;
      uint64_t* d_term = nullptr;
      CU(cudaMalloc((void**)&set->d_rid[i], (n ? n : 1) * 4));
      CU(cudaMallocAsync((void**)&d_term, (ns ? ns : 1) * 8, st));
      CU(cudaMemsetAsync(set->d_rid[i], 0xFF, (n ? n : 1) * 4, st));
      if (ns) {
        int per_sm = 0, sms = 148;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rid_walk, RS_THREADS, 0);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, set->device);
        k_rid_walk<<<grid_for(ns, RS_THREADS, sms * std::max(per_sm, 1)), RS_THREADS, 0, st>>>(v, max_read_len, set->d_rid[i], d_term);
        k_rid_resolve<<<grid_for(n, 256, 148 * 16), 256, 0, st>>>(set->d_rid[i], n, d_term, ix->d_sai, ns);
        if (!getenv("RS_B200_NO_RID_FILL")) /* (tests compare the table with and without pass 3) */
          k_rid_fill<<<grid_for(n, 256, 148 * 8), 256, 0, st>>>(v, ix->d_sai, max_read_len, set->d_rid[i], nullptr);
      }
      CU(cudaGetLastError());
      CU(cudaFreeAsync(d_term, st));
      set->table_bytes += n * 4;
      if (set->have_sets) {
        CU(cudaMalloc((void**)&set->d_sets[i], (ns ? ns : 1) * sizeof(uint4)));
        if (ns) k_read_sets<<<grid_for(ns, 256, 148 * 16), 256, 0, st>>>(ix->d_samp_off, ix->d_samp_codes, (uint32_t)set->sos, ns, set->d_sets[i]);
        CU(cudaGetLastError());
        set->table_bytes += ns * sizeof(uint4);
      }
    }
  } else if (locate_every > 1) {
    for (rs_index* ix : set->shards) {
      if (ix->dev.loc_every != locate_every) {
        const int rc = rs_index_build_locate_samples(ix, max_read_len, locate_every);
        if (rc != RS_STATUS_OK) return rc;
      }
    }
  } else {
    for (rs_index* ix : set->shards) { /* plain walks: a table an earlier seal left on a shard must not answer */
      const int rc = rs_index_drop_locate_table(ix);
      if (rc != RS_STATUS_OK) return rc;
    }
  }

  /* device pointer tables */
  rs_pop_tables& T = set->tab;
  memset(&T, 0, sizeof(T));
  T.nshards = (uint32_t)S;
  T.max_steps = max_read_len;
  T.sos = (uint32_t)set->sos;
  const size_t bytes = S * (sizeof(void*) * 5 + sizeof(uint64_t)) + RS_POP_MAX_SHARD_ID * sizeof(uint32_t) + S * sizeof(rs_index_dev) + 64;
  CU(cudaMalloc(&set->d_tabmem, bytes));
  std::vector<uint8_t> host(bytes, 0);
  size_t o = 0;
  auto put = [&](const void* src, size_t nbytes, size_t align) {
    o = (o + align - 1) / align * align;
    memcpy(host.data() + o, src, nbytes);
    const size_t at = o;
    o += nbytes;
    return (uint8_t*)set->d_tabmem + at;
  };
  std::vector<const void*> p_rid(S), p_sets(S), p_sai(S), p_so(S), p_sc(S);
  std::vector<uint64_t> tag(S);
  std::vector<uint32_t> slot(RS_POP_MAX_SHARD_ID, 0);
  std::vector<rs_index_dev> views(S);
  for (size_t i = 0; i < S; ++i) {
    rs_index* ix = set->shards[i];
    p_rid[i] = set->d_rid[i];
    p_sets[i] = set->d_sets[i];
    p_sai[i] = ix->d_sai;
    p_so[i] = set->have_samples ? ix->d_samp_off : nullptr;
    p_sc[i] = set->have_samples ? ix->d_samp_codes : nullptr;
    tag[i] = (uint64_t)set->shard_ids[i] << 40;
    slot[set->shard_ids[i]] = (uint32_t)i;
    views[i] = ix->dev;
    views[i].loc = views[i].loc_fin = views[i].loc_mark = views[i].loc_rank = nullptr;
    views[i].loc_every = 0;
  }
  T.rid = (const uint32_t* const*)put(p_rid.data(), S * sizeof(void*), 16);
  T.sets = (const uint4* const*)put(p_sets.data(), S * sizeof(void*), 16);
  T.sai = (const uint64_t* const*)put(p_sai.data(), S * sizeof(void*), 16);
  T.samp_off = (const uint64_t* const*)put(p_so.data(), S * sizeof(void*), 16);
  T.samp_codes = (const uint8_t* const*)put(p_sc.data(), S * sizeof(void*), 16);
  T.tag = (const uint64_t*)put(tag.data(), S * sizeof(uint64_t), 16);
  T.slot_of_shard = (const uint32_t*)put(slot.data(), slot.size() * sizeof(uint32_t), 16);
  T.views = (const rs_index_dev*)put(views.data(), S * sizeof(rs_index_dev), 16);
  CU(cudaMemcpyAsync(set->d_tabmem, host.data(), bytes, cudaMemcpyHostToDevice, st));
  CU(cudaStreamSynchronize(st));
  /* Gather passes: the per-read sample-set tables are read at random (one 16-byte set per hit); when a rank's tables
   * exceed what L2 keeps (configs[2] on one GPU: 177 MB over 64 shards) the gather runs over a few shards at a time so
   * that a pass's tables stay L2-resident (RS_POP_PASS_BYTES; 8 shards = 22 MB there).  Power-of-two shard counts keep
   * the lanes of a group on a shard each. */
  set->pass_shards = 0;
  if (set->have_sets && S > 8) {
    uint64_t set_bytes = 0;
    for (rs_index* ix : set->shards) set_bytes += ix->dev.num_strings * sizeof(uint4);
    uint64_t budget = RS_POP_PASS_BYTES;
    if (const char* e = getenv("RS_B200_POP_PASS_BYTES")) budget = (uint64_t)atoll(e); /* 0: one pass */
    if (budget && set_bytes > budget) {
      uint32_t ps = 32;
      while (ps > 8 && set_bytes * ps / S > budget) ps >>= 1;
      if (ps < S) set->pass_shards = ps;
    }
  }
  set->gather_lanes = rs_gather_lanes(set->pass_shards ? set->pass_shards : (uint32_t)set->shards.size());
  set->sort_batches = true;
  if (const char* e = getenv("RS_B200_POP_SORT")) set->sort_batches = atoi(e) != 0;
  set->gather_mode = 0;
  if (const char* e = getenv("RS_B200_GATHER_MODE")) set->gather_mode = !strcmp(e, "shard") ? 2 : !strcmp(e, "query") ? 1 : 0;
  if (const char* e = getenv("RS_B200_GATHER_LANES")) { /* experiments and tests: 1, 2, 4, 8, 16 or 32 */
    const long v = atol(e);
    if (v >= 1 && v <= 32 && (v & (v - 1)) == 0) set->gather_lanes = (uint32_t)v;
  }
  set->sealed = true;
  return RS_STATUS_OK;
}

int rs_shardset_get_info(const rs_shardset* set, rs_shardset_info* info) {
  if (!set || !info) return fail(RS_STATUS_ARG, "null argument");
  memset(info, 0, sizeof(*info));
  info->num_shards = (uint32_t)set->shards.size();
  info->sealed = set->sealed ? 1 : 0;
  info->locate_every = set->locate_every;
  info->max_read_len = set->max_read_len;
  info->num_samples = set->num_samples;
  info->carrier_words = set->words;
  info->table_bytes = set->table_bytes;
  for (rs_index* ix : set->shards) {
    info->index_bytes += ix->hbm_bytes;
    info->bwlen += ix->dev.n;
    info->num_strings += ix->dev.num_strings;
    info->num_runs += ix->num_runs;
  }
  info->chunk_queries = set->chunk_forced ? set->chunk : std::max<uint64_t>(1u << 16, (RS_POP_SCRATCH_BYTES / 12) / std::max<size_t>(1, set->shards.size()));
  return RS_STATUS_OK;
}

int rs_shardset_enable_stats(rs_shardset* set, int on) {
  if (!set) return fail(RS_STATUS_ARG, "null argument");
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  if (on && !set->d_stats) {
    CU(cudaMalloc((void**)&set->d_stats, 32));
    CU(cudaMemset(set->d_stats, 0, 32));
  } else if (!on && set->d_stats) {
    CU(cudaDeviceSynchronize());
    cudaFree(set->d_stats);
    set->d_stats = nullptr;
  }
  return RS_STATUS_OK;
}

int rs_shardset_read_stats(rs_shardset* set, uint64_t* interval_steps, uint64_t* walk_steps, uint64_t* block_loads, int reset) {
  if (!set) return fail(RS_STATUS_ARG, "null argument");
  if (!set->d_stats) return fail(RS_STATUS_NO_TABLE, "statistics are not enabled");
  CU(cudaSetDevice(set->device));
  CU(cudaDeviceSynchronize());
  uint64_t v[4] = {0, 0, 0, 0};
  CU(cudaMemcpy(v, set->d_stats, 32, cudaMemcpyDeviceToHost));
  if (interval_steps) *interval_steps = v[0];
  if (walk_steps) *walk_steps = v[1];
  if (block_loads) *block_loads = v[2];
  if (reset) CU(cudaMemset(set->d_stats, 0, 32));
  return RS_STATUS_OK;
}

int rs_shardset_enable_timing(rs_shardset* set, int on) {
  if (!set) return fail(RS_STATUS_ARG, "null argument");
  std::lock_guard<std::mutex> lk(set->mu);
  set->timing = on != 0;
  set->spans.clear();
  set->ev_used = 0;
  for (int i = 0; i < RS_POP_STAGES; ++i) set->launches[i] = 0;
  return RS_STATUS_OK;
}

int rs_shardset_read_timing(rs_shardset* set, double ms[RS_POP_STAGES], uint64_t launches[RS_POP_STAGES]) {
  if (!set || !ms) return fail(RS_STATUS_ARG, "null argument");
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  CU(cudaDeviceSynchronize());
  for (int i = 0; i < RS_POP_STAGES; ++i) ms[i] = 0.0;
  for (const auto& sp : set->spans) {
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, set->ev_pool[sp.a], set->ev_pool[sp.b]));
    ms[sp.stage] += t;
  }
  if (launches)
    for (int i = 0; i < RS_POP_STAGES; ++i) launches[i] = set->launches[i];
  set->spans.clear();
  set->ev_used = 0;
  for (int i = 0; i < RS_POP_STAGES; ++i) set->launches[i] = 0;
  return RS_STATUS_OK;
}

/* ---- several GPUs: one process per GPU, results fanned in to the query's owner through peer memory ---- */
#define RS_POPCOMM_HDR 256
struct rs_popcomm {
  int device = 0;
  uint32_t rank = 0, world = 1;
  uint64_t maxq = 0, cap = 0;
  uint8_t* base = nullptr;
  uint8_t* peer_base[RS_MAX_RANKS] = {};
  uint64_t** d_peer_flags = nullptr;
  uint64_t epoch = 0;
  bool connected = false, ipc = false;
  size_t o_counts = 0, o_rankh = 0, o_slot = 0, o_off = 0, o_carr = 0, o_total = 0, o_hits = 0, bytes = 0;
  rs_pop_peers peers;
};

int rs_popcomm_create(int device, int rank, int world, uint64_t max_queries, uint64_t hit_capacity, rs_popcomm** out) {
  if (!out || rank < 0 || world < 1 || rank >= world || world > RS_MAX_RANKS || max_queries == 0)
    return fail(RS_STATUS_ARG, "bad argument");
  *out = nullptr;
  CU(cudaSetDevice(device));
  rs_popcomm* c = new rs_popcomm;
  c->device = device;
  c->rank = (uint32_t)rank;
  c->world = (uint32_t)world;
  c->maxq = max_queries;
  c->cap = hit_capacity;
  size_t o = RS_POPCOMM_HDR;
  auto take = [&](size_t nbytes) {
    const size_t at = o;
    o = (o + nbytes + 255) / 256 * 256;
    return at;
  };
  c->o_counts = take(max_queries * 8);
  c->o_rankh = take((size_t)world * max_queries * 4);
  c->o_slot = take(max_queries * 8);
  c->o_off = take((max_queries + 1) * 8);
  c->o_carr = take(max_queries * 16);
  c->o_total = take(8);
  c->o_hits = take((hit_capacity ? hit_capacity : 1) * 8);
  c->bytes = o;
  cudaError_t e = cudaMalloc((void**)&c->base, c->bytes);
  if (e == cudaSuccess) e = cudaMemset(c->base, 0, c->o_hits); /* everything but the hit list */
  if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_peer_flags, RS_MAX_RANKS * sizeof(uint64_t*));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (c->base) cudaFree(c->base);
    delete c;
    return fail(RS_STATUS_CUDA, "rs_popcomm_create: %s", cudaGetErrorString(e));
  }
  c->peer_base[rank] = c->base;
  *out = c;
  return RS_STATUS_OK;
}

int rs_popcomm_export(rs_popcomm* c, uint8_t handle[RS_COMM_HANDLE_BYTES]) {
  if (!c || !handle) return fail(RS_STATUS_ARG, "null argument");
  CU(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, c->base));
  memcpy(handle, &h, sizeof(h));
  return RS_STATUS_OK;
}

static int comm_finish_connect(rs_popcomm* c);

int rs_popcomm_connect(rs_popcomm* c, const uint8_t* handles) {
  if (!c || (!handles && c->world > 1)) return fail(RS_STATUS_ARG, "null argument");
  if (c->connected) return fail(RS_STATUS_ARG, "already connected");
  CU(cudaSetDevice(c->device));
  for (uint32_t r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)r * RS_COMM_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_base[r] = (uint8_t*)p;
  }
  c->ipc = true;
  return comm_finish_connect(c);
}

static int comm_finish_connect(rs_popcomm* c) {
  CU(cudaSetDevice(c->device));
  uint64_t* flags[RS_MAX_RANKS] = {};
  memset(&c->peers, 0, sizeof(c->peers));
  c->peers.world = c->world;
  c->peers.rank = c->rank;
  c->peers.maxq = c->maxq;
  for (uint32_t r = 0; r < c->world; ++r) {
    uint8_t* b = c->peer_base[r];
    flags[r] = reinterpret_cast<uint64_t*>(b);
    c->peers.counts[r] = reinterpret_cast<uint64_t*>(b + c->o_counts);
    c->peers.rankh[r] = reinterpret_cast<uint32_t*>(b + c->o_rankh);
    c->peers.slot[r] = reinterpret_cast<uint64_t*>(b + c->o_slot);
    c->peers.hits[r] = reinterpret_cast<uint64_t*>(b + c->o_hits);
    c->peers.carr[r] = reinterpret_cast<uint32_t*>(b + c->o_carr);
  }
  CU(cudaMemcpy(c->d_peer_flags, flags, sizeof(flags), cudaMemcpyHostToDevice));
  c->connected = true;
  return RS_STATUS_OK;
}

void rs_popcomm_destroy(rs_popcomm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (uint32_t r = 0; r < c->world; ++r)
    if (c->ipc && r != c->rank && c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
  if (c->d_peer_flags) cudaFree(c->d_peer_flags);
  if (c->base) cudaFree(c->base);
  delete c;
}

int rs_popcomm_results(rs_popcomm* c, void** d_counts, void** d_hit_offsets, void** d_hits, void** d_carriers,
                       void** d_total_hits, uint64_t* hit_capacity) {
  if (!c) return fail(RS_STATUS_ARG, "null argument");
  if (d_counts) *d_counts = c->base + c->o_counts;
  if (d_hit_offsets) *d_hit_offsets = c->base + c->o_off;
  if (d_hits) *d_hits = c->base + c->o_hits;
  if (d_carriers) *d_carriers = c->base + c->o_carr;
  if (d_total_hits) *d_total_hits = c->base + c->o_total;
  if (hit_capacity) *hit_capacity = c->cap;
  return RS_STATUS_OK;
}

int rs_popcomm_error(rs_popcomm* c, int* timed_out) {
  if (!c || !timed_out) return fail(RS_STATUS_ARG, "null argument");
  CU(cudaSetDevice(c->device));
  uint32_t v = 0;
  CU(cudaMemcpy(&v, c->base + RS_MAX_RANKS * sizeof(uint64_t), sizeof(v), cudaMemcpyDeviceToHost));
  *timed_out = (int)v;
  return RS_STATUS_OK;
}

static int comm_barrier(rs_popcomm* c, cudaStream_t st) {
  ++c->epoch;
  uint32_t* err = reinterpret_cast<uint32_t*>(c->base + RS_MAX_RANKS * sizeof(uint64_t));
  CU(rs_launch_peer_barrier(c->d_peer_flags, c->world, c->rank, c->epoch, 5000, err, st));
  return RS_STATUS_OK;
}

/* Host-pointer calls: the batch's answers become final range by range of the CALLER's order, so that the copy-out of a
 * range runs while the gather works on the next one.  The batch is then sorted by (range, k-mer): K1 and the offsets
 * see one batch, the shard-major gather runs range after range (its rows still ascend within a range). */
#define RS_POP_MAX_RANGES 4
struct PopProgress {
  uint32_t ranges = 1;                 /* wanted; pop_run stores what it used */
  uint64_t per = 0;                    /* queries per range (pop_run) */
  cudaEvent_t ev_marks = nullptr;      /* recorded once h_marks is on its way */
  cudaEvent_t ev_range[RS_POP_MAX_RANGES] = {};
  uint64_t* h_marks = nullptr;         /* pinned, ranges + 1: hit offsets at the range boundaries */
  uint8_t* d_h40 = nullptr;            /* the list in 5 bytes per hit as well (k_pop_pack40), or null */
  uint32_t* d_bad = nullptr;
};

/* the pipeline proper; everything on `st`, no host synchronisation */
static int pop_run(rs_shardset* set, rs_popcomm* comm, const void* d_kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                   uint64_t* d_counts, uint64_t* d_hit_offsets, uint64_t* d_hits, uint64_t capacity, uint32_t* d_carriers,
                   uint64_t* d_total, cudaStream_t st, bool keep_base = false, PopProgress* prog = nullptr) {
  const uint32_t S = (uint32_t)set->shards.size();
  /* chunk: as many queries as the per-shard interval scratch (12 bytes per shard and query) may take; long launches
   * keep every lane busy (the batch is handed out dynamically) and a shard's index hot in L2 */
  uint64_t C = std::min<uint64_t>(set->chunk, Q);
  if (!set->chunk_forced) {
    const uint64_t fit = std::max<uint64_t>(1u << 16, (RS_POP_SCRATCH_BYTES / 12) / S);
    const uint64_t nchunks = (Q + fit - 1) / fit;
    C = (Q + nchunks - 1) / nchunks; /* equal chunks */
  }
  const bool want_hits = d_hit_offsets != nullptr;
  const bool table = set->locate_every == 1;
  const bool multi = comm != nullptr && comm->world > 1;
  const uint64_t* keys = (const uint64_t*)d_kmers;
  if (!packed) {
    CU(set->s_packed.need(Q * 8));
    StageTimer tm(set, st, RS_POP_STAGE_PACK, 1);
    CU(rs_launch_pack_kmers((const uint8_t*)d_kmers, (uint32_t)k, Q, strand, (uint64_t*)set->s_packed.p, st));
    keys = (const uint64_t*)set->s_packed.p;
  } else if (strand) {
    CU(set->s_packed.need(Q * 8));
    StageTimer tm(set, st, RS_POP_STAGE_PACK, 1);
    k_pop_revcomp<<<(unsigned)((Q + 255) / 256), 256, 0, st>>>(keys, (uint32_t)k, Q, (uint64_t*)set->s_packed.p);
    CU(cudaGetLastError());
    keys = (const uint64_t*)set->s_packed.p;
  }
  CU(set->s_lo.need((size_t)S * C * 8));
  CU(set->s_cnt.need((size_t)S * C * 4));
  CU(set->s_nh.need((C + 1) * 8));
  CU(set->s_loc.need((C + 1) * 8));
  CU(set->s_base.need(8));
  size_t scan_bytes = 0;
  CU(rs_scan_bytes(C, &scan_bytes));
  CU(set->s_scan.need(scan_bytes ? scan_bytes : 8));
  /* large chunks on one GPU: the gather goes shard by shard over the sorted batch; small ones: one launch, query-major.
   * Across GPUs the query-major kernel stays: it writes a query's hits of this rank as one contiguous piece of the
   * owner's list, where the shard-major kernel sends every 8-byte hit over NVLink on its own (8 GPUs, configs[2]:
   * 7.5 ms against 570 ms for the stage); RS_B200_GATHER_MODE=shard still forces it (tests). */
  const bool shard_major = table && want_hits && capacity && (!d_carriers || set->have_sets) &&
                           (set->gather_mode == 2 || (set->gather_mode == 0 && !multi && C >= RS_POP_SHARD_MAJOR_MIN));
  if (want_hits && (!table || shard_major)) CU(set->s_seg.need((size_t)S * C * 4));
  /* Sorted batch (a shard-major gather over the whole batch in one chunk): the packed k-mers are sorted -- first base
   * most significant, which is the order of their intervals in every shard's BWT -- and searched in that order, so the
   * gather of a shard reads its read-index table in ascending row order (streaming instead of one DRAM line per hit)
   * and neighbouring lanes of K1 end in neighbouring blocks.  Answers go to the queries' places in the caller's order
   * through the permutation the sort carries along. */
  const bool sorted = shard_major && set->sort_batches && Q <= C && Q > 1 && Q < (1ull << 32);
  if (shard_major && d_carriers && (multi || sorted)) CU(set->s_carrl.need(C * 16));
  const uint32_t* perm = nullptr;
  const uint64_t* ksearch = keys; /* what K1 and the gather index by position in their launch */
  /* ranges of the caller's order that become final one after the other (host-pointer calls, PopProgress) */
  const uint32_t R = (prog && sorted && Q >= RS_POP_MAX_RANGES) ? std::min<uint32_t>(std::max<uint32_t>(prog->ranges, 1u), RS_POP_MAX_RANGES) : 1u;
  const uint64_t per = (Q + R - 1) / R;
  if (sorted) {
    size_t sort_bytes = 0;
    CU(rs_sort_pairs_bytes(Q, &sort_bytes));
    CU(set->s_sort.need(sort_bytes ? sort_bytes : 8));
    CU(set->s_skeys.need(Q * 8));
    CU(set->s_perm.need(Q * 4));
    CU(set->s_iota.need(Q * 4));
    CU(set->s_sslot.need(Q * 8));
    StageTimer tm(set, st, RS_POP_STAGE_PACK, 1); /* launches counted are this library's own kernels (not CUB's) */
    const unsigned gQ = (unsigned)((Q + 255) / 256);
    k_pop_iota<<<gQ, 256, 0, st>>>((uint32_t*)set->s_iota.p, Q);
    CU(cudaGetLastError());
    if (R > 1) { /* (range, k-mer) order */
      CU(set->s_ckeys.need(Q * 16));
      uint64_t* ck = (uint64_t*)set->s_ckeys.p;
      k_pop_range_keys<<<gQ, 256, 0, st>>>(keys, Q, per, 2u * (uint32_t)k, ck);
      CU(rs_sort_pairs(ck, ck + Q, (const uint32_t*)set->s_iota.p, (uint32_t*)set->s_perm.p, Q, 2 * k + 2, set->s_sort.p,
                       sort_bytes, st));
      k_pop_keys_sorted<<<gQ, 256, 0, st>>>(keys, (const uint32_t*)set->s_perm.p, Q, (uint64_t*)set->s_skeys.p);
      CU(cudaGetLastError());
    } else {
      CU(rs_sort_pairs(keys, (uint64_t*)set->s_skeys.p, (const uint32_t*)set->s_iota.p, (uint32_t*)set->s_perm.p, Q, 2 * k,
                       set->s_sort.p, sort_bytes, st));
    }
    perm = (const uint32_t*)set->s_perm.p;
    ksearch = (const uint64_t*)set->s_skeys.p;
  }
  if (prog) {
    prog->ranges = R;
    prog->per = per;
  }
  const uint32_t pass_shards = (table && !shard_major && want_hits && d_carriers && set->pass_shards) ? set->pass_shards : 0;
  const uint32_t npass = pass_shards ? (S + pass_shards - 1) / pass_shards : 1;
  if (pass_shards) CU(set->s_pass.need((size_t)npass * C * 4));
  uint32_t* pass_off = pass_shards ? (uint32_t*)set->s_pass.p : nullptr;
  uint64_t* lo = (uint64_t*)set->s_lo.p;
  uint32_t* cnt = (uint32_t*)set->s_cnt.p;
  uint64_t* nh = (uint64_t*)set->s_nh.p;
  uint64_t* loc = (uint64_t*)set->s_loc.p;
  uint64_t* base = (uint64_t*)set->s_base.p;
  uint32_t* seg = (want_hits && (!table || shard_major)) ? (uint32_t*)set->s_seg.p : nullptr;
  /* where results go: own buffers (one GPU) or every rank's, through peer memory */
  rs_pop_peers P;
  if (multi) {
    P = comm->peers;
  } else {
    memset(&P, 0, sizeof(P));
    P.world = 1;
    P.maxq = Q;
    P.counts[0] = d_counts;
    P.hits[0] = d_hits;
    P.carr[0] = d_carriers;
    P.slot[0] = d_hit_offsets; /* a query's hits start at its CSR offset */
  }
  const int want_carr = d_carriers != nullptr;
  if (!keep_base) CU(cudaMemsetAsync(base, 0, 8, st)); /* keep_base: this call continues the hit list of the previous one */
  if (multi) {
    /* own accumulation buffers cleared, then a barrier: no peer adds into them earlier */
    CU(cudaMemsetAsync(d_counts, 0, Q * 8, st));
    if (want_carr) CU(cudaMemsetAsync(d_carriers, 0, Q * 16, st));
    const int rc = comm_barrier(comm, st);
    if (rc != RS_STATUS_OK) return rc;
  }
  for (uint64_t q0 = 0; q0 < Q; q0 += C) {
    const uint64_t nq = std::min<uint64_t>(C, Q - q0);
    const unsigned gq = (unsigned)((nq + 255) / 256);
    {
      unsigned long long* tickets = nullptr; /* one work counter per launch, cleared together */
      CU(rs_ticket_take(S, &tickets, st));
      StageTimer tm(set, st, RS_POP_STAGE_SEARCH, S);
      for (uint32_t s = 0; s < S; ++s)
        CU(rs_launch_find_interval_packed(set->shards[s]->dev, ksearch + q0, (uint32_t)k, nq, lo + (uint64_t)s * C,
                                          cnt + (uint64_t)s * C, 4, (unsigned long long*)set->d_stats, st, nullptr, tickets + s));
    }
    {
      StageTimer tm(set, st, RS_POP_STAGE_OFFSETS, (want_hits ? 3 : 1) + (multi ? (want_hits ? 5 : 1) : 0));
      if (!multi) {
        k_pop_totals<<<gq, 256, 0, st>>>(cnt, S, C, nq, max_hits, d_counts + q0, nh, seg, pass_off, pass_shards, perm);
      } else {
        if (seg) k_pop_totals<<<gq, 256, 0, st>>>(cnt, S, C, nq, max_hits, loc /* unused sums */, nh, seg, nullptr, 0, nullptr);
        k_pop_totals_comm<<<gq, 256, 0, st>>>(P, ksearch + q0, (uint32_t)k, cnt, S, C, q0, nq, max_hits, want_hits ? 1 : 0, pass_off,
                                              pass_shards, perm);
      }
      CU(cudaGetLastError());
      if (!want_hits) continue;
      if (multi) {
        int rc = comm_barrier(comm, st); /* every rank's hit numbers of this chunk have arrived */
        if (rc != RS_STATUS_OK) return rc;
        k_pop_owner_hits<<<gq, 256, 0, st>>>(P, keys + q0, (uint32_t)k, q0, nq, nh);
      }
      CU(rs_exclusive_scan(nh, loc, nq, set->s_scan.p, scan_bytes, st));
      k_pop_add_base<<<(unsigned)((nq + 256) / 256), 256, 0, st>>>(loc, nq, base, d_hit_offsets + q0);
      k_pop_advance<<<1, 1, 0, st>>>(loc, nq, base, q0 + nq == Q ? d_hit_offsets + Q : nullptr, q0 + nq == Q ? d_total : nullptr);
      CU(cudaGetLastError());
      if (prog && R > 1) { /* the list's extent per range, on its way to the host while the gather runs */
        for (uint32_t r = 0; r <= R; ++r)
          CU(cudaMemcpyAsync(&prog->h_marks[r], d_hit_offsets + std::min<uint64_t>((uint64_t)r * per, Q), 8, cudaMemcpyDeviceToHost, st));
        CU(cudaEventRecord(prog->ev_marks, st));
      }
      if (multi) {
        k_pop_owner_slots<<<gq, 256, 0, st>>>(P, keys + q0, (uint32_t)k, q0, nq, d_hit_offsets);
        CU(cudaGetLastError());
        const int rc = comm_barrier(comm, st); /* every owner's slots have arrived */
        if (rc != RS_STATUS_OK) return rc;
      }
    }
    if (!capacity) { /* sizing call: offsets and total only (carrier sets need reported hits) */
      if (d_carriers && !multi) CU(cudaMemsetAsync(d_carriers + q0 * set->words, 0, nq * (size_t)set->words * 4, st));
      continue;
    }
    const uint64_t* slot = P.slot[multi ? comm->rank : 0];
    StageTimer tm_locate(set, st, RS_POP_STAGE_LOCATE, shard_major ? S + 3 : table ? npass : 5 * S + 1);
    if (shard_major) {
      uint4* carr = nullptr;
      if (want_carr) { /* the chunk's sets start empty: the launches OR into them */
        carr = (multi || sorted) ? (uint4*)set->s_carrl.p : reinterpret_cast<uint4*>(d_carriers + q0 * 4);
        CU(cudaMemsetAsync(carr, 0, nq * 16, st));
      }
      if (sorted) { /* q0 = 0: where the hits of the i-th query in sorted order start */
        k_pop_slot_sorted<<<gq, 256, 0, st>>>(slot, perm, nq, (uint64_t*)set->s_sslot.p);
        slot = (const uint64_t*)set->s_sslot.p;
      }
      if (R > 1) { /* sorted by (range, k-mer), q0 = 0: range after range, each final when its last launch ends */
        for (uint32_t r = 0; r < R; ++r) {
          const uint64_t a = std::min<uint64_t>((uint64_t)r * per, Q), b = std::min<uint64_t>(a + per, Q), n = b - a;
          if (n) {
            const int gs = grid_for(n, 256, 148 * RS_GATHER_BLOCKS);
            for (uint32_t sh = 0; sh < S; ++sh)
              k_pop_gather_shard<<<gs, 256, 0, st>>>(set->tab, P, sh, ksearch + a, (uint32_t)k, lo + (uint64_t)sh * C + a,
                                                     cnt + (uint64_t)sh * C + a, seg + (uint64_t)sh * C + a, 0, n, max_hits,
                                                     slot + a, capacity, carr ? carr + a : nullptr);
            if (carr) k_pop_carr_unsort<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(carr + a, perm + a, n, reinterpret_cast<uint4*>(d_carriers));
            if (prog->d_h40) k_pop_pack40<<<148 * 4, 256, 0, st>>>(d_hits, d_hit_offsets + a, d_hit_offsets + b, capacity, prog->d_h40, prog->d_bad);
            CU(cudaGetLastError());
          }
          CU(cudaEventRecord(prog->ev_range[r], st));
        }
      } else {
        const int gs = grid_for(nq, 256, 148 * RS_GATHER_BLOCKS);
        for (uint32_t sh = 0; sh < S; ++sh)
          k_pop_gather_shard<<<gs, 256, 0, st>>>(set->tab, P, sh, ksearch + q0, (uint32_t)k, lo + (uint64_t)sh * C, cnt + (uint64_t)sh * C,
                                                 seg + (uint64_t)sh * C, q0, nq, max_hits, slot, capacity, carr);
        if (carr && multi) k_pop_carr_send<<<gq, 256, 0, st>>>(P, ksearch + q0, (uint32_t)k, q0, nq, carr, perm);
        else if (carr && sorted) k_pop_carr_unsort<<<gq, 256, 0, st>>>(carr, perm, nq, reinterpret_cast<uint4*>(d_carriers));
        CU(cudaGetLastError());
      }
    } else if (table) {
      const int wc = (want_carr && set->have_sets) ? 1 : 0;
      const int gg = 148 * 2 * RS_GATHER_BLOCKS;
#define RS_GATHER(G)                                                                                                       \
  k_pop_gather<G><<<grid_for(nq * G, 256, gg), 256, 0, st>>>(set->tab, P, keys + q0, (uint32_t)k, lo, cnt, C, q0, nq, max_hits, \
                                                             slot, capacity, wc, sb, se, pass_off ? pass_off + (size_t)ps * C : nullptr)
      for (uint32_t ps = 0; ps < npass; ++ps) {
        const uint32_t sb = pass_shards ? ps * pass_shards : 0, se = pass_shards ? std::min<uint32_t>(S, sb + pass_shards) : S;
        switch (set->gather_lanes) {
          case 32: RS_GATHER(32); break;
          case 16: RS_GATHER(16); break;
          case 8: RS_GATHER(8); break;
          case 4: RS_GATHER(4); break;
          case 2: RS_GATHER(2); break;
          default: RS_GATHER(1); break;
        }
      }
#undef RS_GATHER
      CU(cudaGetLastError());
      if (want_carr && !set->have_sets) {
        k_pop_carriers<<<grid_for(nq * 32, 256, 148 * 8), 256, 8 * set->words * 4, st>>>(set->tab, d_hit_offsets + q0, nq, d_hits, capacity,
                                                                                      set->words, d_carriers + q0 * set->words);
        CU(cudaGetLastError());
      }
    } else {
      /* shards too large for the read-index table: per shard, rows -> walk (K2 / sampled) -> scatter */
      CU(set->s_lcnt.need((C + 1) * 8));
      uint64_t* lcnt = (uint64_t*)set->s_lcnt.p;
      uint64_t* loff = nh; /* nh is free again after the scan above */
      /* a chunk's hits in one shard are bounded by the caller's capacity as well */
      const uint64_t bound = std::min<uint64_t>(capacity, nq * std::min<uint64_t>(max_hits, 0xFFFFFFFFull));
      CU(set->s_rows.need((bound ? bound : 1) * 8));
      CU(set->s_dst.need((bound ? bound : 1) * 8));
      CU(set->s_ri.need((bound ? bound : 1) * 8));
      for (uint32_t s = 0; s < S; ++s) {
        rs_index* ix = set->shards[s];
        k_pop_shard_counts<<<gq, 256, 0, st>>>(cnt + (uint64_t)s * C, nq, max_hits, lcnt);
        CU(rs_exclusive_scan(lcnt, loff, nq, set->s_scan.p, scan_bytes, st));
        k_pop_expand<<<grid_for(nq * 32, 256, 148 * 8), 256, 0, st>>>(P, keys + q0, (uint32_t)k, lo + (uint64_t)s * C, loff, q0, nq, slot,
                                                                   seg + (uint64_t)s * C, bound, capacity,
                                                                   (uint64_t*)set->s_rows.p, (uint64_t*)set->s_dst.p);
        CU(cudaGetLastError());
        CU(rs_launch_extract_prefix_n(ix->dev, (const uint64_t*)set->s_rows.p, loff + nq, bound, set->max_read_len, ix->d_sai,
                                      (uint64_t*)set->s_ri.p, (unsigned long long*)set->d_stats, st));
        k_pop_scatter<<<148 * 8, 256, 0, st>>>((const uint64_t*)set->s_ri.p, (const uint64_t*)set->s_dst.p, loff + nq, bound,
                                              (uint64_t)set->shard_ids[s] << 40);
        CU(cudaGetLastError());
      }
      if (want_carr) {
        k_pop_carriers<<<grid_for(nq * 32, 256, 148 * 8), 256, 8 * set->words * 4, st>>>(set->tab, d_hit_offsets + q0, nq, d_hits, capacity,
                                                                                      set->words, d_carriers + q0 * set->words);
        CU(cudaGetLastError());
      }
    }
  }
  if (!want_hits && d_total) CU(cudaMemsetAsync(d_total, 0, 8, st));
  if (multi) {
    const int rc = comm_barrier(comm, st); /* every rank's adds, hits and bitmaps have landed */
    if (rc != RS_STATUS_OK) return rc;
  }
  if (prog && R == 1) { /* one range: everything is final here */
    if (want_hits) {
      if (prog->d_h40 && capacity) {
        k_pop_pack40<<<148 * 8, 256, 0, st>>>(d_hits, d_hit_offsets, d_hit_offsets + Q, capacity, prog->d_h40, prog->d_bad);
        CU(cudaGetLastError());
      }
      CU(cudaMemcpyAsync(&prog->h_marks[0], d_hit_offsets, 8, cudaMemcpyDeviceToHost, st));
      CU(cudaMemcpyAsync(&prog->h_marks[1], d_hit_offsets + Q, 8, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaEventRecord(prog->ev_marks, st));
    CU(cudaEventRecord(prog->ev_range[0], st));
  }
  return RS_STATUS_OK;
}

/* Several GPUs: every rank calls this with the same batch on its own shard set; the answers of the k-mers a rank owns
 * (first four bases modulo the ranks) end up in that rank's communicator buffers (rs_popcomm_results). */
int rs_pop_query_comm_dev(rs_shardset* set, rs_popcomm* comm, const void* d_kmers, int k, size_t Q, int packed, int strand,
                          uint64_t max_hits, int want_hits, int want_carriers, void* stream) {
  if (!set || !comm || (Q && !d_kmers)) return fail(RS_STATUS_ARG, "null argument");
  if (!set->sealed) return fail(RS_STATUS_ARG, "shard set is not sealed");
  if (!comm->connected) return fail(RS_STATUS_ARG, "communicator not connected");
  if (set->device != comm->device) return fail(RS_STATUS_ARG, "shard set and communicator live on different devices");
  if (k < 1 || k > 31) return fail(RS_STATUS_ARG, "population queries are fixed-length k-mers of 1..31 characters");
  if (Q > comm->maxq) return fail(RS_STATUS_ARG, "batch larger than the communicator's capacity");
  if (want_carriers && !want_hits) return fail(RS_STATUS_ARG, "carrier sets need the hit list");
  if (want_carriers && comm->world > 1 && !set->have_sets)
    return fail(RS_STATUS_NO_TABLE, "carrier sets across GPUs need the per-read sample-set tables (<= 128 samples, locate_every = 1)");
  if (want_carriers && !set->have_samples) return fail(RS_STATUS_NO_TABLE, "carrier sets requested but a shard has no sample table");
  if (want_carriers && set->words != 4) return fail(RS_STATUS_TOO_LARGE, "the communicator holds 128-sample carrier sets");
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  uint8_t* b = comm->base;
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 (forward) or 1 (reverse complement)");
  return pop_run(set, comm, d_kmers, k, Q, packed, strand, max_hits, (uint64_t*)(b + comm->o_counts),
                 want_hits ? (uint64_t*)(b + comm->o_off) : nullptr, (uint64_t*)(b + comm->o_hits), want_hits ? comm->cap : 0,
                 want_carriers ? (uint32_t*)(b + comm->o_carr) : nullptr, (uint64_t*)(b + comm->o_total), (cudaStream_t)stream);
}

static int pop_check(rs_shardset* set, const void* kmers, int k, size_t Q, int strand, const void* counts,
                     const void* hit_offsets, const void* hits, uint64_t capacity, const void* carriers) {
  if (!set || (Q && (!kmers || !counts))) return fail(RS_STATUS_ARG, "null argument");
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 (forward) or 1 (reverse complement)");
  if (!set->sealed) return fail(RS_STATUS_ARG, "shard set is not sealed");
  if (k < 1 || k > 31) return fail(RS_STATUS_ARG, "population queries are fixed-length k-mers of 1..31 characters");
  if (hit_offsets && capacity && !hits) return fail(RS_STATUS_ARG, "capacity given but hits is null");
  if (carriers && !hit_offsets) return fail(RS_STATUS_ARG, "carrier sets need the hit list (hit_offsets / hits)");
  if (carriers && !set->have_samples) return fail(RS_STATUS_NO_TABLE, "carrier sets requested but a shard has no sample table");
  return RS_STATUS_OK;
}

int rs_pop_query_dev(rs_shardset* set, const void* d_kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                     void* d_counts, void* d_hit_offsets, void* d_hits, uint64_t capacity, void* d_carriers,
                     void* d_total_hits, void* stream) {
  const int rc = pop_check(set, d_kmers, k, Q, strand, d_counts, d_hit_offsets, d_hits, capacity, d_carriers);
  if (rc != RS_STATUS_OK) return rc;
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  if (!Q) {
    if (d_hit_offsets) CU(cudaMemsetAsync(d_hit_offsets, 0, 8, (cudaStream_t)stream));
    if (d_total_hits) CU(cudaMemsetAsync(d_total_hits, 0, 8, (cudaStream_t)stream));
    return RS_STATUS_OK;
  }
  return pop_run(set, nullptr, d_kmers, k, Q, packed, strand, max_hits, (uint64_t*)d_counts, (uint64_t*)d_hit_offsets, (uint64_t*)d_hits,
                 capacity, (uint32_t*)d_carriers, (uint64_t*)d_total_hits, (cudaStream_t)stream);
}

/* host buffers (pinned or not): the batch goes through the pipeline in a few sub-batches so that the copy-out of one
 * (its slice of counts, offsets, carrier sets and -- the bulk -- its hits) overlaps the kernels of the next;
 * RS_STATUS_CAPACITY when hits were dropped */
static int pop_query_host(rs_shardset* set, const void* kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                          uint64_t* counts, uint64_t* hit_offsets, void* hits_out, int hit_bytes, uint64_t capacity,
                          uint32_t* carriers, uint64_t* total_hits) {
  uint64_t* hits = static_cast<uint64_t*>(hits_out);       /* hit_bytes == 8 */
  uint8_t* hits40 = static_cast<uint8_t*>(hits_out);       /* hit_bytes == 5 */
  const int rc = pop_check(set, kmers, k, Q, strand, counts, hit_offsets, hits, capacity, carriers);
  if (rc != RS_STATUS_OK) return rc;
  if (total_hits) *total_hits = 0;
  if (!Q) {
    if (hit_offsets) hit_offsets[0] = 0;
    return RS_STATUS_OK;
  }
  std::lock_guard<std::mutex> lk(set->mu);
  CU(cudaSetDevice(set->device));
  cudaStream_t st = set->stream;
  if (!set->copy_stream) {
    CU(cudaStreamCreateWithFlags(&set->copy_stream, cudaStreamNonBlocking));
    CU(cudaMallocHost((void**)&set->h_marks, RS_POP_MAX_SUB * sizeof(uint64_t)));
    for (int i = 0; i < RS_POP_MAX_SUB; ++i) CU(cudaEventCreateWithFlags(&set->ev_sub[i], cudaEventDisableTiming));
  }
  cudaStream_t cs = set->copy_stream;
  const size_t unit = packed ? 8 : (size_t)k;
  const size_t W = set->words;
  CU(set->s_in.need(Q * unit));
  CU(set->s_hc.need(Q * 8));
  CU(set->s_ho.need((Q + 1) * 8));
  CU(set->s_hm.need(RS_POP_MAX_SUB * 8));
  if (hit_offsets && capacity) CU(set->s_hh.need(capacity * 8));
  uint8_t* d_h40 = nullptr;
  uint32_t* d_bad = nullptr;
  if (hit_bytes == 5 && hit_offsets && capacity) {
    CU(set->s_h40.need(capacity * 5 + 64));
    d_h40 = (uint8_t*)set->s_h40.p + 8;                     /* the flag word, then the 8-byte aligned list */
    d_bad = (uint32_t*)set->s_h40.p;
    CU(cudaMemsetAsync(d_bad, 0, 8, st));
  }
  if (carriers) CU(set->s_hw.need(Q * W * 4));
  uint64_t *d_counts = (uint64_t*)set->s_hc.p, *d_off = hit_offsets ? (uint64_t*)set->s_ho.p : nullptr;
  uint64_t *d_hits = (hit_offsets && capacity) ? (uint64_t*)set->s_hh.p : nullptr, *d_marks = (uint64_t*)set->s_hm.p;
  uint32_t* d_carr = carriers ? (uint32_t*)set->s_hw.p : nullptr;
  /* sub-batches of up to 16 M queries (they bound the device-side result buffers of very large batches), at most
   * RS_POP_MAX_SUB.  Smaller ones, whose copy-out overlaps the next one's kernels, measured slower on configs[2]: every
   * sub-batch streams all shards' rank blocks and tables through L2 again (10 M 31-mers, 5-byte hits: 1 sub-batch
   * 43.7 M q/s, 4 sub-batches 39.7 M q/s) */
  size_t sub_queries = (size_t)1 << 24;
  if (const char* e = getenv("RS_B200_POP_SUB")) sub_queries = (size_t)std::max<long long>(1000, atoll(e)); /* experiments */
  size_t nsub = (Q + sub_queries - 1) / sub_queries;
  if (nsub > RS_POP_MAX_SUB) nsub = RS_POP_MAX_SUB;
  if (!hit_offsets) nsub = 1;
  const size_t per = (Q + nsub - 1) / nsub;
  CU(cudaMemcpyAsync(set->s_in.p, kmers, Q * unit, cudaMemcpyHostToDevice, st));
  /* Copy-out range by range of the caller's order: pop_run sorts a large batch by (range, k-mer) and gathers range after
   * range, so a range's answers leave while the next one is gathered (RS_B200_POP_RANGES: 1..4, experiments) */
  uint32_t want_ranges = (hit_offsets && capacity) ? RS_POP_MAX_RANGES : 1;
  if (const char* e = getenv("RS_B200_POP_RANGES")) want_ranges = (uint32_t)std::min<long>(RS_POP_MAX_RANGES, std::max<long>(1, atol(e)));
  uint64_t total = 0;
  int status = RS_STATUS_OK;
  for (size_t b = 0; b < nsub && status == RS_STATUS_OK; ++b) {
    const size_t q0 = b * per, nq = std::min(per, Q - q0);
    PopProgress pg;
    pg.ranges = want_ranges;
    pg.ev_marks = set->ev_sub[RS_POP_MAX_RANGES];
    for (int r = 0; r < RS_POP_MAX_RANGES; ++r) pg.ev_range[r] = set->ev_sub[r];
    pg.h_marks = set->h_marks;
    pg.d_h40 = d_h40;
    pg.d_bad = d_bad;
    status = pop_run(set, nullptr, (const uint8_t*)set->s_in.p + q0 * unit, k, nq, packed, strand, max_hits, d_counts + q0,
                     d_off ? d_off + q0 : nullptr, d_hits, capacity, d_carr ? d_carr + q0 * W : nullptr, d_marks + b, st, b > 0,
                     &pg);
    if (status != RS_STATUS_OK) break;
    CU(cudaEventSynchronize(pg.ev_marks)); /* the ranges' extents in the hit list are on the host */
    for (uint32_t r = 0; r < pg.ranges; ++r) {
      const size_t ra = std::min<size_t>((size_t)r * pg.per, nq), re = std::min<size_t>(ra + pg.per, nq);
      CU(cudaStreamWaitEvent(cs, pg.ev_range[r], 0));
      if (re == ra) continue;
      CU(cudaMemcpyAsync(counts + q0 + ra, d_counts + q0 + ra, (re - ra) * 8, cudaMemcpyDeviceToHost, cs));
      if (!hit_offsets) continue;
      CU(cudaMemcpyAsync(hit_offsets + q0 + ra, d_off + q0 + ra, (re - ra + (q0 + re == Q ? 1 : 0)) * 8, cudaMemcpyDeviceToHost, cs));
      if (carriers) CU(cudaMemcpyAsync(carriers + (q0 + ra) * W, d_carr + (q0 + ra) * W, (re - ra) * W * 4, cudaMemcpyDeviceToHost, cs));
      const uint64_t hb = std::min<uint64_t>(set->h_marks[r], capacity), he = std::min<uint64_t>(set->h_marks[r + 1], capacity);
      if (hits_out && he > hb) {
        if (d_h40) CU(cudaMemcpyAsync(hits40 + hb * 5, d_h40 + hb * 5, (he - hb) * 5, cudaMemcpyDeviceToHost, cs));
        else CU(cudaMemcpyAsync(hits + hb, d_hits + hb, (he - hb) * 8, cudaMemcpyDeviceToHost, cs));
      }
    }
    if (hit_offsets) total = set->h_marks[pg.ranges];
  }
  cudaStreamSynchronize(st);
  const cudaError_t ce = cudaStreamSynchronize(cs);
  if (status != RS_STATUS_OK) return status;
  if (ce != cudaSuccess) return fail(RS_STATUS_CUDA, "copy-out: %s", cudaGetErrorString(ce));
  if (total_hits) *total_hits = total;
  if (hit_offsets && total > capacity)
    return fail(RS_STATUS_CAPACITY, "%llu hits do not fit the %llu provided", (unsigned long long)total,
                (unsigned long long)capacity);
  if (d_bad) {
    uint32_t bad = 0;
    CU(cudaMemcpy(&bad, d_bad, 4, cudaMemcpyDeviceToHost));
    if (bad) return fail(RS_STATUS_TOO_LARGE, "a hit does not fit 40 bits (shard id >= 255 or read index >= 2^32): use rs_pop_query");
  }
  return RS_STATUS_OK;
}

int rs_pop_query(rs_shardset* set, const void* kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                 uint64_t* counts, uint64_t* hit_offsets, uint64_t* hits, uint64_t capacity, uint32_t* carriers,
                 uint64_t* total_hits) {
  return pop_query_host(set, kmers, k, Q, packed, strand, max_hits, counts, hit_offsets, hits, 8, capacity, carriers, total_hits);
}

int rs_pop_query_hits40(rs_shardset* set, const void* kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                        uint64_t* counts, uint64_t* hit_offsets, uint8_t* hits40, uint64_t capacity, uint32_t* carriers,
                        uint64_t* total_hits) {
  return pop_query_host(set, kmers, k, Q, packed, strand, max_hits, counts, hit_offsets, hits40, 5, capacity, carriers, total_hits);
}
