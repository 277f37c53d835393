/* K4 on the device: SGA run bytes -> 64-byte blocks + position / select directories (rs_layout.h), and the
 * constants of the reference's tail behaviour.  Takes the place of RLEBWT::initialiseFMIndex
 * (src/bwt/rlebwt.cpp:34-148), which walks the runs once on one host thread; the same structure is built here by
 * a handful of data-parallel passes, because a block is a FIXED number of runs (36):
 *
 *   k_enc_sums      per block: symbols per character in its 36 runs (+ validation of every run byte)
 *   CUB scans       exclusive prefix sums over blocks -> symbols / occurrences before each block
 *   k_enc_blocks    header + payload of every block
 *   k_enc_dir       position directory: block b fills the 512-symbol slots whose first symbol it holds
 *   k_enc_sel       select directories: block b fills the entries whose occurrence it holds
 *   k_enc_tail      one thread: replay of the m_vSum hint rule (rlebwt.cpp:103-106) and of the bucket arithmetic of
 *                   BPTree::rank / access (BPTree.h:74-94, 135-157) -> clamp positions and initInterval's constants
 *
 * rs_build.cpp (host, one streaming pass) builds the same arrays; tests compare the two byte for byte.
 */
#include "rs_internal.h"

#include <cub/device/device_scan.cuh>

#include "rs_encode.h"

/* per block: [0] symbols, [1..4] occurrences of A,C,G,T ('$' = symbols - sum) -- u64 each for the scans */
__global__ void k_enc_sums(const uint8_t* __restrict__ runs, uint64_t R, uint64_t nblocks, uint64_t* __restrict__ sums,
                           unsigned int* __restrict__ bad) {
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblocks) return;
  const uint64_t r0 = b * RS_PAYLOAD;
  uint32_t tot = 0, c[5] = {0, 0, 0, 0, 0};
  for (uint32_t t = 0; t < RS_PAYLOAD && r0 + t < R; ++t) {
    const uint32_t v = runs[r0 + t], sym = v >> 5, len = v & 31u;
    if (sym > 4 || len == 0) {
      atomicOr(bad, 1u);
      continue;
    }
    c[sym] += len;
    tot += len;
  }
  sums[b] = tot;
  sums[(nblocks + 1) + b] = c[1];
  sums[2 * (nblocks + 1) + b] = c[2];
  sums[3 * (nblocks + 1) + b] = c[3];
  sums[4 * (nblocks + 1) + b] = c[4];
}

__global__ void k_enc_blocks(const uint8_t* __restrict__ runs, uint64_t R, uint64_t nblocks,
                             const uint64_t* __restrict__ before, rs_block* __restrict__ blocks,
                             unsigned int* __restrict__ bad) {
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblocks) return;
  rs_block k;
  k.start = before[b];
  k.cnt_hi = 0;
  for (int c = 0; c < 4; ++c) {
    const uint64_t v = before[(uint64_t)(c + 1) * (nblocks + 1) + b];
    if (v >> 40) atomicOr(bad, 2u);
    k.cnt_lo[c] = (uint32_t)v;
    k.cnt_hi |= (uint32_t)((v >> 32) & 0xFFu) << (8 * c);
  }
  const uint64_t r0 = b * RS_PAYLOAD;
  for (uint32_t t = 0; t < RS_PAYLOAD; ++t) k.runs[t] = r0 + t < R ? runs[r0 + t] : 0;
  blocks[b] = k;
}

/* start of block q, or n for q == nblocks */
__device__ __forceinline__ uint64_t enc_start(const uint64_t* __restrict__ before, uint64_t nblocks, uint64_t q) {
  return before[q]; /* before[nblocks] = n: the scans run over nblocks + 1 entries */
}

__global__ void k_enc_dir(const uint64_t* __restrict__ before, uint64_t nblocks, uint64_t ndir, uint64_t* __restrict__ dir) {
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblocks) return;
  const uint64_t s0 = before[b], s1 = before[b + 1];
  /* slots whose first symbol p = j << 9 lies in [s0, s1); the last block also takes every slot past the end */
  uint64_t j0 = (s0 + RS_SLOT_MASK) >> RS_SLOT_SHIFT;
  uint64_t j1 = b + 1 == nblocks ? ndir : (s1 + RS_SLOT_MASK) >> RS_SLOT_SHIFT;
  if (b == 0) j0 = 0;
  for (uint64_t j = j0; j < j1 && j < ndir; ++j) {
    const uint64_t p = j << RS_SLOT_SHIFT;
    uint64_t e = b;
    int nb = 0;
    for (uint64_t q = b + 1; q < nblocks && before[q] < p + RS_SLOT_SIZE; ++q) {
      if (nb < 3) e |= (before[q] - p) << (32 + RS_SLOT_SHIFT * nb);
      ++nb;
    }
    if (nb > 3) e |= RS_DIR_MORE;
    dir[j] = e;
  }
}

/* entry k of symbol c = the block with occ_before[c] <= k << sh < occ_before_next[c] */
__global__ void k_enc_sel(const uint64_t* __restrict__ before, uint64_t nblocks, uint32_t c, uint32_t sh, uint64_t nent,
                          uint32_t* __restrict__ sel) {
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblocks) return;
  uint64_t lo, hi;
  if (c == 0) {
    lo = before[b];
    hi = before[b + 1];
    for (int x = 1; x <= 4; ++x) {
      lo -= before[(uint64_t)x * (nblocks + 1) + b];
      hi -= before[(uint64_t)x * (nblocks + 1) + b + 1];
    }
  } else {
    lo = before[(uint64_t)c * (nblocks + 1) + b];
    hi = before[(uint64_t)c * (nblocks + 1) + b + 1];
  }
  const uint64_t one = 1ull << sh;
  for (uint64_t k = (lo + one - 1) >> sh; (k << sh) < hi && k < nent; ++k) sel[k] = (uint32_t)b;
}

__global__ void k_fill_u32(uint32_t* __restrict__ p, uint64_t n, uint32_t v) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

/* symbols (out[0]) and occurrences of $,A,C,G,T (out[1..5]) before run number r (r <= R) */
__device__ void enc_before_run(const uint8_t* __restrict__ runs, const uint64_t* __restrict__ before, uint64_t nblocks,
                               uint64_t r, uint64_t* out) {
  const uint64_t b = r / RS_PAYLOAD; /* r == R with R a multiple of 36 gives b == nblocks: the totals */
  uint64_t c[5];
  out[0] = before[b];
  c[0] = before[b];
  for (int x = 1; x <= 4; ++x) {
    c[x] = before[(uint64_t)x * (nblocks + 1) + b];
    c[0] -= c[x];
  }
  for (uint64_t t = b * RS_PAYLOAD; t < r; ++t) {
    const uint32_t v = runs[t];
    c[v >> 5] += v & 31u;
    out[0] += v & 31u;
  }
  for (int x = 0; x < 5; ++x) out[1 + x] = c[x];
}

__global__ void k_enc_tail(const uint8_t* __restrict__ runs, uint64_t R, const uint64_t* __restrict__ before, uint64_t nblocks,
                           uint64_t depth, rs_enc_tail* __restrict__ out) {
  if (threadIdx.x || blockIdx.x) return;
  const uint64_t n = before[nblocks];
  uint64_t tot[6];
  enc_before_run(runs, before, nblocks, R, tot);
  for (int x = 0; x < 5; ++x) out->tot[x] = tot[1 + x];
  out->n = n;
  /* m_vSum: a hint is pushed at the first run whose start has reached the next multiple of 65,536
   * (rlebwt.cpp:103-106); run lengths are <= 31, so one hint per multiple: the K-th at run 1 + (run holding
   * symbol K * 65536 - 1) */
  const uint64_t last_start = n - (runs[R - 1] & 31u);
  const uint64_t K = last_start >> 16;
  uint64_t vsum_last = 0;
  if (K) {
    const uint64_t target = (K << 16) - 1; /* symbol position */
    uint64_t lo = 0, hi = nblocks - 1;     /* last block with start <= target */
    while (lo < hi) {
      const uint64_t mid = (lo + hi + 1) >> 1;
      if (before[mid] <= target) lo = mid;
      else hi = mid - 1;
    }
    uint64_t pos = before[lo], r = lo * RS_PAYLOAD;
    while (pos + (runs[r] & 31u) <= target) pos += runs[r++] & 31u; /* run r holds the target symbol */
    vsum_last = ((r + 1) - 1) >> 6;
  }
  out->vlast = K;
  const uint64_t shift0 = (depth > 1 ? 4 : 0) + (depth > 2 ? 10 * (depth - 2) : 0);
  const uint64_t top_runs = 64ull << shift0;
  const uint64_t next_run = ((vsum_last >> shift0) + 1) * top_runs;
  uint64_t at[6];
  if (next_run < R) enc_before_run(runs, before, nblocks, next_run, at);
  else for (int x = 0; x < 6; ++x) at[x] = tot[x];
  out->clamp_c = at[0];
  /* BPTree::access keeps descending by hint while the bucket index on the hint's path is 0 (BPTree.h:148-151) */
  uint64_t j = 0, sv = shift0;
  while (j + 1 < depth) {
    if ((vsum_last >> sv) != 0) break;
    ++j;
    const uint64_t rem = depth - j;
    sv = (rem > 1 ? 4 : 0) + (rem > 2 ? 10 * (rem - 2) : 0);
  }
  const uint64_t next_run_char = ((vsum_last >> sv) + 1) * (64ull << sv);
  if (next_run_char == next_run) out->clamp_char = out->clamp_c;
  else if (next_run_char >= R) out->clamp_char = n;
  else {
    uint64_t t[6];
    enc_before_run(runs, before, nblocks, next_run_char, t);
    out->clamp_char = t[0];
  }
  const bool clamped = (n >> 16) >= K && n - 1 >= out->clamp_c;
  for (int x = 0; x < 5; ++x) out->init_occ[x] = clamped ? at[1 + x] : tot[1 + x];
  out->q2_safe = (n >> 16) <= K ? 1 : 0;
}

/* Encodes d_runs (device, R bytes) into freshly allocated device arrays.  One host synchronisation (the totals
 * size the select directories). */
int rs_encode_device(const uint8_t* d_runs, uint64_t R, uint64_t num_symbols, cudaStream_t st, rs_encoded* out) {
  memset(out, 0, sizeof(*out));
  if (!d_runs || R == 0) return RS_ERR_ARG;
  const uint64_t nblocks = (R + RS_PAYLOAD - 1) / RS_PAYLOAD;
  if (nblocks > 0xFFFFFFFFull) return RS_ERR_TOO_LARGE;
  uint64_t depth = 1;
  for (uint64_t x = R >> 10; x > 0; x >>= 10) ++depth;
  out->depth = depth;
  out->nblocks = nblocks;
  uint64_t *d_sums = nullptr, *d_before = nullptr;
  unsigned int* d_bad = nullptr;
  rs_enc_tail* d_tail = nullptr;
  void* d_tmp = nullptr;
  size_t tmp_bytes = 0;
  int rc = RS_OK;
  cudaError_t e = cudaSuccess;
  const size_t col = nblocks + 1;
#define ENC_CU(x)                 \
  do {                            \
    e = (x);                      \
    if (e != cudaSuccess) goto done; \
  } while (0)
  {
    ENC_CU(cudaMalloc((void**)&d_sums, 5 * col * 8));
    ENC_CU(cudaMalloc((void**)&d_before, 5 * col * 8));
    ENC_CU(cudaMalloc((void**)&d_bad, 4));
    ENC_CU(cudaMalloc((void**)&d_tail, sizeof(rs_enc_tail)));
    ENC_CU(cudaMemsetAsync(d_bad, 0, 4, st));
    ENC_CU(cudaMemsetAsync(d_sums, 0, 5 * col * 8, st));
    const unsigned gb = (unsigned)((nblocks + 255) / 256);
    k_enc_sums<<<gb, 256, 0, st>>>(d_runs, R, nblocks, d_sums, d_bad);
    ENC_CU(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_sums, d_before, col, st));
    ENC_CU(cudaMalloc(&d_tmp, tmp_bytes ? tmp_bytes : 8));
    for (int c = 0; c < 5; ++c)
      ENC_CU(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_sums + c * col, d_before + c * col, col, st));
    ENC_CU(cudaMalloc(&out->d_blocks, nblocks * sizeof(rs_block)));
    k_enc_blocks<<<gb, 256, 0, st>>>(d_runs, R, nblocks, d_before, (rs_block*)out->d_blocks, d_bad);
    k_enc_tail<<<1, 1, 0, st>>>(d_runs, R, d_before, nblocks, depth, d_tail);
    ENC_CU(cudaGetLastError());
    unsigned int bad = 0;
    ENC_CU(cudaMemcpyAsync(&out->tail, d_tail, sizeof(rs_enc_tail), cudaMemcpyDeviceToHost, st));
    ENC_CU(cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, st));
    ENC_CU(cudaStreamSynchronize(st));
    if (bad & 1u) {
      rc = RS_ERR_RUNBYTE;
      goto done;
    }
    if (bad & 2u) {
      rc = RS_ERR_TOO_LARGE;
      goto done;
    }
    if (out->tail.n != num_symbols) {
      rc = RS_ERR_ARG;
      goto done;
    }
    const uint64_t n = out->tail.n;
    out->ndir = (n >> RS_SLOT_SHIFT) + 2;
    ENC_CU(cudaMalloc(&out->d_dir, out->ndir * 8));
    k_enc_dir<<<gb, 256, 0, st>>>(d_before, nblocks, out->ndir, (uint64_t*)out->d_dir);
    for (int c = 0; c < RS_NSYM; ++c) {
      uint32_t sh = 0;
      while (sh < RS_SEL_SHIFT_MAX && (out->tail.tot[c] >> (sh + 1)) >= nblocks / 2 + 1) ++sh;
      out->sel_shift[c] = sh;
      out->nsel[c] = (out->tail.tot[c] >> sh) + 2;
      ENC_CU(cudaMalloc(&out->d_sel[c], out->nsel[c] * 4));
      k_fill_u32<<<(unsigned)((out->nsel[c] + 255) / 256), 256, 0, st>>>((uint32_t*)out->d_sel[c], out->nsel[c], (uint32_t)(nblocks - 1));
      k_enc_sel<<<gb, 256, 0, st>>>(d_before, nblocks, (uint32_t)c, sh, out->nsel[c], (uint32_t*)out->d_sel[c]);
    }
    ENC_CU(cudaGetLastError());
    ENC_CU(cudaStreamSynchronize(st));
  }
done:
#undef ENC_CU
  if (d_sums) cudaFree(d_sums);
  if (d_before) cudaFree(d_before);
  if (d_bad) cudaFree(d_bad);
  if (d_tail) cudaFree(d_tail);
  if (d_tmp) cudaFree(d_tmp);
  if (e != cudaSuccess) {
    rs_fail(RS_STATUS_CUDA, "device encoder: %s", cudaGetErrorString(e));
    rc = RS_ERR_CUDA;
  }
  if (rc != RS_OK) rs_encoded_free(out);
  return rc;
}

void rs_encoded_free(rs_encoded* e) {
  if (e->d_blocks) cudaFree(e->d_blocks);
  if (e->d_dir) cudaFree(e->d_dir);
  for (int c = 0; c < RS_NSYM; ++c)
    if (e->d_sel[c]) cudaFree(e->d_sel[c]);
  e->d_blocks = e->d_dir = nullptr;
  for (int c = 0; c < RS_NSYM; ++c) e->d_sel[c] = nullptr;
}
