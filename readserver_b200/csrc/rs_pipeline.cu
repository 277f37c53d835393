/* Glue kernels between the FM-index kernels: interval -> hit rows, read index -> sample codes,
 * hits -> per-k-mer carrier sets.  Streaming integer work; prefix sums via CUB (toolkit header).
 *
 *   hit expansion   the per-row loop of query()            src/bwt/query.cpp:93-97
 *   sample gather   RocksDB Get + value decode             src/service/service.cpp:816-845
 */
#include "rs_kernels.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

/* counts[q] = rows of interval q, capped at max_hits (count_reads' count, service.cpp:304) */
__global__ void k_hit_counts(const uint64_t* __restrict__ lower, const uint64_t* __restrict__ upper, uint64_t Q,
                             uint64_t max_hits, uint64_t* __restrict__ counts) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  const uint64_t lo = lower[q], up = upper[q];
  uint64_t c = up >= lo ? up - lo + 1 : 0;
  counts[q] = c < max_hits ? c : max_hits;
}

/* rows[offsets[q] + j] = lower[q] + j; one warp per interval, coalesced stores */
__global__ void k_expand_rows(const uint64_t* __restrict__ lower, const uint64_t* __restrict__ offsets, uint64_t Q,
                              uint64_t capacity, uint64_t* __restrict__ rows, uint32_t* __restrict__ owner) {
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  for (uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < Q; q += warps) {
    const uint64_t b = offsets[q], e = offsets[q + 1], lo = lower[q];
    for (uint64_t j = b + lane; j < e && j < capacity; j += 32) {
      rows[j] = lo + (j - b);
      if (owner) owner[j] = (uint32_t)q;
    }
  }
}

cudaError_t rs_scan_bytes(uint64_t n, size_t* bytes) {
  *bytes = 0;
  return cub::DeviceScan::ExclusiveSum(nullptr, *bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, n + 1);
}

/* offsets[0..n] = exclusive prefix sum of counts[0..n) (counts[n] is ignored: must be allocated) */
cudaError_t rs_exclusive_scan(const uint64_t* counts, uint64_t* offsets, uint64_t n, void* tmp, size_t tmp_bytes,
                              cudaStream_t st) {
  return cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, offsets, n + 1, st);
}

/* (key, value) pairs sorted by bits [0, end_bit) of the key: the shard set sorts a batch's packed k-mers (first base
 * most significant = the order of their BWT intervals) so that the locate stage reads each read-index table in
 * ascending row order */
cudaError_t rs_sort_pairs_bytes(uint64_t n, size_t* bytes) {
  *bytes = 0;
  return cub::DeviceRadixSort::SortPairs(nullptr, *bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                         (const uint32_t*)nullptr, (uint32_t*)nullptr, n, 0, 64);
}
cudaError_t rs_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                          uint64_t n, int end_bit, void* tmp, size_t tmp_bytes, cudaStream_t st) {
  return cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, st);
}

cudaError_t rs_launch_hit_counts(const uint64_t* lower, const uint64_t* upper, uint64_t Q, uint64_t max_hits,
                                 uint64_t* counts, cudaStream_t st) {
  if (!Q) return cudaSuccess;
  k_hit_counts<<<(unsigned)((Q + 255) / 256), 256, 0, st>>>(lower, upper, Q, max_hits, counts);
  return cudaGetLastError();
}

cudaError_t rs_launch_expand_rows(const uint64_t* lower, const uint64_t* offsets, uint64_t Q, uint64_t capacity,
                                  uint64_t* rows, uint32_t* owner, cudaStream_t st) {
  if (!Q) return cudaSuccess;
  uint64_t blocks = (Q * 32 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_expand_rows<<<(unsigned)blocks, 256, 0, st>>>(lower, offsets, Q, capacity, rows, owner);
  return cudaGetLastError();
}

/* ---- samples ---- */
/* counts[i] = codes of read read_index[i] (0 for UINT64_MAX / out-of-range reads) */
__global__ void k_sample_counts(const uint64_t* __restrict__ samp_off, uint64_t num_strings,
                                const uint64_t* __restrict__ read_index, uint64_t n, uint64_t* __restrict__ counts) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t r = read_index[i];
  counts[i] = r < num_strings ? samp_off[r + 1] - samp_off[r] : 0;
}

/* out_codes[out_off[i] ..] = codes of read read_index[i]; one warp per read */
__global__ void k_sample_gather(const uint64_t* __restrict__ samp_off, const uint8_t* __restrict__ samp_codes,
                                uint32_t sos, uint64_t num_strings, const uint64_t* __restrict__ read_index, uint64_t n,
                                const uint64_t* __restrict__ out_off, uint8_t* __restrict__ out_codes) {
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
    const uint64_t r = read_index[i];
    if (r >= num_strings) continue;
    const uint64_t src = samp_off[r] * sos, bytes = (samp_off[r + 1] - samp_off[r]) * sos, dst = out_off[i] * sos;
    for (uint64_t j = lane; j < bytes; j += 32) out_codes[dst + j] = samp_codes[src + j];
  }
}

cudaError_t rs_launch_sample_counts(const uint64_t* samp_off, uint64_t num_strings, const uint64_t* read_index,
                                    uint64_t n, uint64_t* counts, cudaStream_t st) {
  if (!n) return cudaSuccess;
  k_sample_counts<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(samp_off, num_strings, read_index, n, counts);
  return cudaGetLastError();
}

cudaError_t rs_launch_sample_gather(const uint64_t* samp_off, const uint8_t* samp_codes, uint32_t sos,
                                    uint64_t num_strings, const uint64_t* read_index, uint64_t n,
                                    const uint64_t* out_off, uint8_t* out_codes, cudaStream_t st) {
  if (!n) return cudaSuccess;
  uint64_t blocks = (n * 32 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_sample_gather<<<(unsigned)blocks, 256, 0, st>>>(samp_off, samp_codes, sos, num_strings, read_index, n, out_off,
                                                   out_codes);
  return cudaGetLastError();
}

/* ---- carrier sets: per k-mer, the sorted set of distinct samples over all its hit reads ---- */
/* sample code (create_samples_hash.cpp:31-51): base-223 digits + 33, most significant first, 1-based */
__device__ __forceinline__ uint32_t rs_code_to_sample(const uint8_t* code, uint32_t sos) {
  uint32_t v = 0;
  for (uint32_t i = 0; i < sos; ++i) v = v * 223u + (uint32_t)(code[i] - 33u);
  return v - 1u;
}

__device__ __forceinline__ void rs_sample_to_code(uint32_t sample, uint32_t sos, uint8_t* out) {
  uint32_t rem = sample + 1u;
  for (uint32_t i = 1; i <= sos; ++i) {
    out[sos - i] = (uint8_t)(rem % 223u + 33u);
    rem /= 223u;
  }
}

/* One warp per k-mer.  A bitmap over samples lives in shared memory (words per warp = bm_words).
 * emit == 0: counts[q] = distinct samples.   emit == 1: write their codes, ascending, at out_off[q]. */
__global__ void k_carriers(const uint64_t* __restrict__ samp_off, const uint8_t* __restrict__ samp_codes, uint32_t sos,
                           uint64_t num_strings, uint32_t num_samples, uint32_t bm_words,
                           const uint64_t* __restrict__ hit_off, const uint64_t* __restrict__ read_index, uint64_t Q,
                           int emit, uint64_t* __restrict__ counts, const uint64_t* __restrict__ out_off,
                           uint8_t* __restrict__ out_codes) {
  extern __shared__ uint32_t smem[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  uint32_t* bm = smem + wib * bm_words;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < Q; q += warps) {
    for (uint32_t w = lane; w < bm_words; w += 32) bm[w] = 0;
    __syncwarp();
    const uint64_t b = hit_off[q], e = hit_off[q + 1];
    for (uint64_t j = b + lane; j < e; j += 32) {
      const uint64_t r = read_index[j];
      if (r >= num_strings) continue;
      for (uint64_t t = samp_off[r]; t < samp_off[r + 1]; ++t) {
        const uint32_t s = rs_code_to_sample(samp_codes + t * sos, sos);
        if (s < num_samples) atomicOr(&bm[s >> 5], 1u << (s & 31));
      }
    }
    __syncwarp();
    /* rank of each word's bits via a warp scan over word popcounts, 32 words at a time */
    uint32_t base = 0;
    for (uint32_t w0 = 0; w0 < bm_words; w0 += 32) {
      const uint32_t w = w0 + lane;
      const uint32_t bits = w < bm_words ? bm[w] : 0u;
      uint32_t pc = __popc(bits), incl = pc;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((int)lane >= d) incl += t;
      }
      if (emit) {
        uint32_t pos = base + incl - pc, rest = bits;
        while (rest) {
          const uint32_t bit = __ffs(rest) - 1;
          rest &= rest - 1;
          rs_sample_to_code(w * 32 + bit, sos, out_codes + (out_off[q] + pos) * sos);
          ++pos;
        }
      }
      base += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    if (!emit && lane == 0) counts[q] = base;
    __syncwarp();
  }
}

cudaError_t rs_launch_carriers(const uint64_t* samp_off, const uint8_t* samp_codes, uint32_t sos, uint64_t num_strings,
                               uint32_t num_samples, const uint64_t* hit_off, const uint64_t* read_index, uint64_t Q,
                               int emit, uint64_t* counts, const uint64_t* out_off, uint8_t* out_codes,
                               cudaStream_t st) {
  if (!Q) return cudaSuccess;
  const uint32_t bm_words = (num_samples + 31) / 32;
  const size_t smem = (size_t)8 * bm_words * sizeof(uint32_t);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  uint64_t blocks = (Q * 32 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_carriers<<<(unsigned)blocks, 256, smem, st>>>(samp_off, samp_codes, sos, num_strings, num_samples, bm_words, hit_off,
                                                  read_index, Q, emit, counts, out_off, out_codes);
  return cudaGetLastError();
}
