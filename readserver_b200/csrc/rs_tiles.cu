/* K-mer tiling of long queries on the device (first slice).
 *
 * /kmermatch cuts a query into tiles of `kmer` bases starting every skip+1 bases (get_tiles,
 * src/service/service.cpp:232-246), drops tiles holding anything but ACGT (find_kmer_reads, :479-481) and
 * searches each tile (find_reads -> findInterval).  Here the tiles of a whole batch of queries are cut by one
 * kernel into a dense [T, kmer] array, searched by K1 in one launch, and tiles the reference skips come back
 * as the empty interval (1, 0).  Duplicate tiles of one query (the reference keeps a set) are searched twice and
 * yield the same interval: the union of their reads is unchanged.
 */
#include "rs_internal.h"

__global__ void k_gather_tiles(const uint8_t* __restrict__ text, const uint64_t* __restrict__ offs,
                               const uint64_t* __restrict__ tile_off, uint64_t Q, uint32_t kmer, uint32_t step,
                               uint8_t* __restrict__ tiles, uint8_t* __restrict__ skipped) {
  /* one warp per query: lane l copies bases l, l+32, ... of each tile */
  const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (q >= Q) return;
  const uint8_t* w = text + offs[q];
  const uint64_t t0 = tile_off[q], nt = tile_off[q + 1] - t0;
  for (uint64_t t = 0; t < nt; ++t) {
    const uint8_t* src = w + t * step;
    uint8_t* dst = tiles + (t0 + t) * kmer;
    bool bad = false;
    for (uint32_t i = lane; i < kmer; i += 32) {
      const uint8_t ch = src[i];
      dst[i] = ch;
      bad |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T');
    }
    bad = __any_sync(0xFFFFFFFFu, bad);
    if (lane == 0) skipped[t0 + t] = bad ? 1 : 0;
  }
}

__global__ void k_void_skipped(const uint8_t* __restrict__ skipped, uint64_t T, uint64_t* __restrict__ lower,
                               uint64_t* __restrict__ upper) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < T && skipped[t]) {
    lower[t] = 1;
    upper[t] = 0;
  }
}

int rs_tile_intervals_batch(const rs_index* ix, const char* text, const uint64_t* offsets, size_t Q, uint32_t kmer,
                            uint32_t skip, uint64_t* tile_offsets, uint64_t capacity, uint64_t* lower, uint64_t* upper) {
  if (!ix || !tile_offsets || (Q && (!text || !offsets))) return fail(RS_STATUS_ARG, "null argument");
  if (kmer < 1) return fail(RS_STATUS_ARG, "kmer must be >= 1");
  const uint64_t step = (uint64_t)skip + 1;
  /* get_tiles: i = 0, skip+1, ... while i <= length - kmer */
  tile_offsets[0] = 0;
  for (size_t q = 0; q < Q; ++q) {
    const uint64_t len = offsets[q + 1] - offsets[q];
    tile_offsets[q + 1] = tile_offsets[q] + (len < kmer ? 0 : (len - kmer) / step + 1);
  }
  const uint64_t T = tile_offsets[Q];
  if (T > capacity)
    return fail(RS_STATUS_CAPACITY, "%llu tiles do not fit the %llu provided", (unsigned long long)T,
                (unsigned long long)capacity);
  if (!T) return RS_STATUS_OK;
  if (!lower || !upper) return fail(RS_STATUS_ARG, "null argument");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  const size_t tbytes = (size_t)offsets[Q];
  CU(ix->s_in.need(tbytes ? tbytes : 1));
  CU(ix->s_in2.need((Q + 1) * 8));
  CU(ix->s_out.need(T * 8));
  CU(ix->s_out2.need(T * 8));
  CU(ix->s_out3.need((Q + 1) * 8));
  CU(ix->s_out4.need(T * (size_t)kmer + T));
  uint8_t* d_tiles = (uint8_t*)ix->s_out4.p;
  uint8_t* d_skipped = d_tiles + T * (size_t)kmer;
  CU(cudaMemcpyAsync(ix->s_in.p, text, tbytes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ix->s_in2.p, offsets, (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ix->s_out3.p, tile_offsets, (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  k_gather_tiles<<<(unsigned)((Q * 32 + 255) / 256), 256, 0, st>>>((const uint8_t*)ix->s_in.p, (const uint64_t*)ix->s_in2.p,
                                                                  (const uint64_t*)ix->s_out3.p, Q, kmer, (uint32_t)step,
                                                                  d_tiles, d_skipped);
  CU(cudaGetLastError());
  CU(rs_launch_find_interval(ix->dev, d_tiles, nullptr, kmer, T, (uint64_t*)ix->s_out.p, (uint64_t*)ix->s_out2.p, 0, 0,
                             nullptr, st));
  k_void_skipped<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(d_skipped, T, (uint64_t*)ix->s_out.p, (uint64_t*)ix->s_out2.p);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(lower, ix->s_out.p, T * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(upper, ix->s_out2.p, T * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}
