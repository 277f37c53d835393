/* Seed stage of /gt on the device (second slice).
 *
 * For a site query the reference cuts the window around the site into tiles and, per tile, collects the reads
 * that hold the tile and reach the site (find_gt_reads, src/service/service.cpp:507-711, one call per tile index
 * from GtTask::run, :1049-1072):
 *
 *   - the tile's interval; while it is wider than max_interval_size (10,000, :91) the tile is grown by one base
 *     and searched again -- to the right, then (a second seed) to the left, or the other way round, depending on
 *     which side of the tile the site lies;
 *   - every row of the final interval(s) is walked both ways; a read is kept when it can reach the site:
 *       tile left of the site   pos - end  > postfix.size() - kmers + 4  -> dropped
 *       tile right of the site  start - pos > prefix.size() + 4          -> dropped
 *       tile over the site      all kept
 *     (all in size_t: after an extension `pos - end` / `start - pos` may wrap, and then every row is dropped);
 *   - the union over the tiles is the read set handed to the aligner.
 *
 * Here: k_gt_seeds computes the (up to) two final intervals of every tile of every query, one thread per tile
 * with the same unsigned arithmetic; the rows of all seeds are expanded and walked by K2 in one launch
 * (prefix length = offset of the match in the read, read index through the .sai); postfix lengths come from the
 * packed read table when one is attached and rank is exact, else from the select walk (k_extract_postfix);
 * k_gt_filter applies the reach test; a radix sort + unique over (query, read index) gives each query's read set.
 * The aligner (ksw) stays on the host.
 *
 * Where the reference leaves defined behaviour the tile yields no seed and the query is flagged:
 *   RS_GT_FLAG_THROW  the extension moved `start` below 1: w.substr(start-1, ..) throws std::out_of_range
 *   RS_GT_FLAG_HANG   the extension ran past the end of w: substr truncates, the same string is searched again
 *                     and the loop never ends
 *   RS_GT_FLAG_WALK   a row's walk did not end within max_steps (tail emulation only)
 */
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>

#include "rs_internal.h"
#include "rs_gt_ops.h"

namespace {

/* one thread per tile: seeds 2t and 2t+1 */
__global__ void k_gt_seeds(const __grid_constant__ rs_index_dev ix, const uint8_t* __restrict__ text,
                           const uint64_t* __restrict__ offs, const uint64_t* __restrict__ tile_off,
                           const uint64_t* __restrict__ pos, uint64_t Q, uint64_t T, uint32_t kmers, uint32_t step,
                           uint64_t max_iv, uint64_t* __restrict__ s_lo, uint64_t* __restrict__ s_hi,
                           uint64_t* __restrict__ s_thr, uint8_t* __restrict__ s_kind, uint32_t* __restrict__ s_q,
                           uint32_t* __restrict__ flags) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  /* query of this tile: last q with tile_off[q] <= t */
  uint64_t a = 0, b = Q;
  while (b - a > 1) {
    const uint64_t m = (a + b) >> 1;
    if (tile_off[m] <= t) a = m;
    else b = m;
  }
  const uint64_t q = a;
  const uint64_t index = t - tile_off[q];
  const uint8_t* w = text + offs[q];
  const uint64_t L = offs[q + 1] - offs[q];
  const uint64_t p = pos[q];
  gt_iv seed[2];
  uint64_t thr[2];
  uint32_t kind, flag;
  rs_gt_tile_seeds(ix, w, L, p, index, kmers, step, max_iv, seed, thr, &kind, &flag);
  for (int i = 0; i < 2; ++i) {
    s_lo[2 * t + i] = seed[i].lo;
    s_hi[2 * t + i] = seed[i].hi;
    s_thr[2 * t + i] = thr[i];
    s_kind[2 * t + i] = (uint8_t)kind;
    s_q[2 * t + i] = (uint32_t)q;
  }
  if (flag) atomicOr(flags + q, flag);
}

/* key[j] = query * num_strings + read index for the rows that reach the site, else `dropped` (above every key) */
__global__ void k_gt_filter(const uint32_t* __restrict__ owner, const uint32_t* __restrict__ pre_steps,
                            const uint64_t* __restrict__ read_index, const uint32_t* __restrict__ post_steps,
                            const uint64_t* __restrict__ read_off, uint64_t H, const uint64_t* __restrict__ s_thr,
                            const uint8_t* __restrict__ s_kind, const uint32_t* __restrict__ s_q, uint32_t kmers,
                            uint64_t num_strings, uint64_t dropped, uint64_t* __restrict__ key,
                            uint32_t* __restrict__ flags) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= H) return;
  const uint32_t s = owner[j], q = s_q[s], kind = s_kind[s];
  const uint32_t pre = pre_steps[j];
  const uint64_t r = read_index[j];
  uint64_t out = dropped;
  if (pre == 0xFFFFFFFFu || r >= num_strings) {
    atomicOr(flags + q, RS_GT_FLAG_WALK);
  } else {
    bool keep = true;
    const uint64_t thr = s_thr[s];
    if (kind == RS_GT_NEED_PREFIX) {
      keep = !(thr > (uint64_t)pre + 4);
    } else if (kind == RS_GT_NEED_POSTFIX) {
      uint64_t post; /* postfix.size(): the match and what follows it */
      if (post_steps) {
        post = post_steps[j];
        if (post_steps[j] == 0xFFFFFFFFu) {
          atomicOr(flags + q, RS_GT_FLAG_WALK);
          keep = false;
        }
      } else {
        post = read_off[r + 1] - read_off[r] - pre;
      }
      keep = keep && !(thr > post - kmers + 4);
    }
    if (keep) out = (uint64_t)q * num_strings + r;
  }
  key[j] = out;
}

/* out_off[q] = first unique key >= q * num_strings, q = 0..Q */
__global__ void k_gt_offsets(const uint64_t* __restrict__ uniq, uint64_t M, uint64_t Q, uint64_t num_strings,
                             uint64_t* __restrict__ out_off) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q > Q) return;
  const uint64_t want = q * num_strings;
  uint64_t a = 0, b = M;
  while (a < b) {
    const uint64_t m = (a + b) >> 1;
    if (uniq[m] < want) a = m + 1;
    else b = m;
  }
  out_off[q] = a;
}

__global__ void k_gt_strip(uint64_t* __restrict__ uniq, uint64_t M, uint64_t num_strings) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < M) uniq[i] %= num_strings;
}

__global__ void k_seed_counts(const uint64_t* __restrict__ lo, const uint64_t* __restrict__ hi, uint64_t S,
                              uint64_t* __restrict__ counts) {
  const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < S) counts[s] = hi[s] >= lo[s] ? hi[s] - lo[s] + 1 : 0;
}

inline unsigned blocks_for(uint64_t n) { return (unsigned)((n + 255) / 256); }

}  // namespace

int rs_gt_reads_batch(const rs_index* ix, const char* text, const uint64_t* offsets, const uint64_t* pos, size_t Q,
                      uint32_t kmer, uint32_t skip, uint64_t max_interval_size, uint32_t max_steps, uint64_t max_rows,
                      uint64_t* out_offsets, uint64_t capacity, uint64_t* read_index, uint32_t* flags,
                      uint64_t* candidate_rows) {
  if (!ix || !out_offsets || !flags || (Q && (!text || !offsets || !pos))) return fail(RS_STATUS_ARG, "null argument");
  if (kmer < 1) return fail(RS_STATUS_ARG, "kmer must be >= 1");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  if (!ix->d_sai) return fail(RS_STATUS_NO_TABLE, "the read sets need the .sai table (rs_index_set_sai)");
  if (Q >= ((uint64_t)1 << 32) || (ix->dev.num_strings && Q > ~0ull / ix->dev.num_strings - 1))
    return fail(RS_STATUS_TOO_LARGE, "too many queries in one batch");
  const uint64_t step = (uint64_t)skip + 1;
  if (step >> 32) return fail(RS_STATUS_ARG, "skip too large");
  /* GtTask::run: tiles_size = (wsize - kmer) / (skip + 1) + 1, nothing when pos > wsize (:1044-1049);
   * a window shorter than one tile (size_t wrap in the reference) has no tiles here */
  std::vector<uint64_t> tile_off(Q + 1, 0);
  for (size_t q = 0; q < Q; ++q) {
    const uint64_t len = offsets[q + 1] - offsets[q];
    const bool none = len < kmer || pos[q] > len;
    tile_off[q + 1] = tile_off[q] + (none ? 0 : (len - kmer) / step + 1);
    flags[q] = 0;
  }
  if (candidate_rows) *candidate_rows = 0;
  const uint64_t T = tile_off[Q], S = 2 * T;
  for (size_t q = 0; q <= Q; ++q) out_offsets[q] = 0;
  if (!T) return RS_STATUS_OK;

  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  Tmp tmp(st);
  const size_t tbytes = (size_t)offsets[Q];
  uint8_t *d_text, *d_kind;
  uint64_t *d_offs, *d_toff, *d_pos, *d_lo, *d_hi, *d_thr, *d_cnt, *d_hoff;
  uint32_t *d_sq, *d_flags;
  void* d_scan;
  size_t scan_bytes = 0;
  CU(tmp.get((void**)&d_text, tbytes));
  CU(tmp.get((void**)&d_offs, (Q + 1) * 8));
  CU(tmp.get((void**)&d_toff, (Q + 1) * 8));
  CU(tmp.get((void**)&d_pos, Q * 8));
  CU(tmp.get((void**)&d_lo, S * 8));
  CU(tmp.get((void**)&d_hi, S * 8));
  CU(tmp.get((void**)&d_thr, S * 8));
  CU(tmp.get((void**)&d_kind, S));
  CU(tmp.get((void**)&d_sq, S * 4));
  CU(tmp.get((void**)&d_flags, Q * 4));
  CU(tmp.get((void**)&d_cnt, (S + 1) * 8));
  CU(tmp.get((void**)&d_hoff, (S + 1) * 8));
  CU(rs_scan_bytes(S, &scan_bytes));
  CU(tmp.get(&d_scan, scan_bytes));
  CU(cudaMemcpyAsync(d_text, text, tbytes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_offs, offsets, (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_toff, tile_off.data(), (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_pos, pos, Q * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(d_flags, 0, Q * 4, st));
  k_gt_seeds<<<blocks_for(T), 256, 0, st>>>(ix->dev, d_text, d_offs, d_toff, d_pos, Q, T, kmer, (uint32_t)step,
                                            max_interval_size, d_lo, d_hi, d_thr, d_kind, d_sq, d_flags);
  CU(cudaGetLastError());
  k_seed_counts<<<blocks_for(S), 256, 0, st>>>(d_lo, d_hi, S, d_cnt);
  CU(cudaGetLastError());
  CU(rs_exclusive_scan(d_cnt, d_hoff, S, d_scan, scan_bytes, st));
  uint64_t H = 0;
  CU(cudaMemcpyAsync(&H, d_hoff + S, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(flags, d_flags, Q * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (candidate_rows) *candidate_rows = H;
  if (H > max_rows)
    return fail(RS_STATUS_CAPACITY, "%llu candidate rows exceed max_rows = %llu", (unsigned long long)H,
                (unsigned long long)max_rows);
  if (!H) return RS_STATUS_OK;

  uint64_t *d_rows, *d_ri, *d_key, *d_sorted, *d_uniq, *d_num, *d_ooff;
  uint32_t *d_owner, *d_pre, *d_post = nullptr;
  CU(tmp.get((void**)&d_rows, H * 8));
  CU(tmp.get((void**)&d_owner, H * 4));
  CU(tmp.get((void**)&d_pre, H * 4));
  CU(tmp.get((void**)&d_ri, H * 8));
  CU(tmp.get((void**)&d_key, H * 8));
  CU(tmp.get((void**)&d_sorted, H * 8));
  CU(tmp.get((void**)&d_uniq, H * 8));
  CU(tmp.get((void**)&d_num, 8));
  CU(tmp.get((void**)&d_ooff, (Q + 1) * 8));
  CU(rs_launch_expand_rows(d_lo, d_hoff, S, H, d_rows, d_owner, st));
  CU(rs_launch_extract_prefix(ix->dev, d_rows, H, max_steps, nullptr, d_pre, ix->d_sai, d_ri, nullptr,
                              (unsigned long long*)ix->d_stats, st));
  const bool from_table = ix->d_read_off && !ix->dev.emulate_tail;
  if (!from_table) {
    CU(tmp.get((void**)&d_post, H * 4));
    CU(rs_launch_extract_postfix(ix->dev, d_rows, H, max_steps, nullptr, d_post, nullptr, st));
  }
  const uint64_t top = (uint64_t)Q * ix->dev.num_strings; /* every kept key is below it */
  k_gt_filter<<<blocks_for(H), 256, 0, st>>>(d_owner, d_pre, d_ri, d_post, ix->d_read_off, H, d_thr, d_kind, d_sq, kmer,
                                             ix->dev.num_strings, top, d_key, d_flags);
  CU(cudaGetLastError());
  /* (query, read) pairs -> sorted unique */
  size_t sort_bytes = 0, uniq_bytes = 0;
  int end_bit = 1;
  while (end_bit < 64 && (top >> end_bit)) ++end_bit;
  CU(cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, d_key, d_sorted, H, 0, end_bit, st));
  CU(cub::DeviceSelect::Unique(nullptr, uniq_bytes, d_sorted, d_uniq, d_num, H, st));
  void* d_work;
  CU(tmp.get(&d_work, sort_bytes > uniq_bytes ? sort_bytes : uniq_bytes));
  CU(cub::DeviceRadixSort::SortKeys(d_work, sort_bytes, d_key, d_sorted, H, 0, end_bit, st));
  CU(cub::DeviceSelect::Unique(d_work, uniq_bytes, d_sorted, d_uniq, d_num, H, st));
  uint64_t M = 0, last = 0;
  CU(cudaMemcpyAsync(&M, d_num, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (M) {
    CU(cudaMemcpyAsync(&last, d_uniq + (M - 1), 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (last == top) --M;
  }
  k_gt_offsets<<<blocks_for(Q + 1), 256, 0, st>>>(d_uniq, M, Q, ix->dev.num_strings, d_ooff);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out_offsets, d_ooff, (Q + 1) * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(flags, d_flags, Q * 4, cudaMemcpyDeviceToHost, st));
  if (M > capacity || (M && !read_index)) {
    CU(cudaStreamSynchronize(st));
    return fail(RS_STATUS_CAPACITY, "%llu reads do not fit the %llu provided", (unsigned long long)M,
                (unsigned long long)capacity);
  }
  if (M) {
    k_gt_strip<<<blocks_for(M), 256, 0, st>>>(d_uniq, M, ix->dev.num_strings);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(read_index, d_uniq, M * 8, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}
