/* Packed read table: output=reads without walking.
 *
 * The reference rebuilds every matching read by walking the BWT both ways from the hit row
 * (query(): extractPrefix + extractPostfix, src/bwt/query.cpp:87-100; find_reads, service.cpp:714-751):
 * about read-length dependent rank/select steps per read.  The locate stage already yields the read's index
 * in builder input order, so the string can be read from a table instead: all reads concatenated at 2 bits
 * per base (reads holding anything but ACGT never reach the BWT, build_bwt.sh:919), addressed by read index.
 * 25 bytes of sequential HBM per 100-base read instead of ~100 random 64-byte blocks.
 */
#include "rs_internal.h"

__global__ void k_pack_reads(const uint8_t* __restrict__ text, uint64_t nbases, uint32_t* __restrict__ packed,
                             unsigned int* __restrict__ bad) {
  const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t b0 = w * 16;
  if (b0 >= nbases) return;
  uint32_t v = 0;
  for (uint32_t i = 0; i < 16 && b0 + i < nbases; ++i) {
    const uint32_t ch = text[b0 + i];
    /* (ch >> 1) & 3: A 0, C 1, T 2, G 3 -> codes A 0, C 1, G 2, T 3 */
    const uint32_t sel = (ch >> 1) & 3u;
    const uint32_t code = (0x2310u >> (4 * sel)) & 3u;
    if (ch != ((0x47544341u >> (8 * sel)) & 0xFFu)) atomicOr(bad, 1u);
    v |= code << (2 * i);
  }
  packed[w] = v;
}

/* one warp per requested read */
__global__ void k_unpack_reads(const uint32_t* __restrict__ packed, const uint64_t* __restrict__ off, uint64_t num_strings,
                               const uint64_t* __restrict__ read_index, uint64_t n, uint32_t max_len,
                               uint8_t* __restrict__ out, uint32_t* __restrict__ out_len) {
  const uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (i >= n) return;
  const uint64_t r = read_index[i];
  uint8_t* dst = out + i * (uint64_t)max_len;
  uint32_t len = 0;
  uint64_t base = 0;
  if (r < num_strings) {
    base = off[r];
    len = (uint32_t)(off[r + 1] - base);
  }
  const uint32_t m = len < max_len ? len : max_len;
  if (((max_len | (uint32_t)(uintptr_t)out) & 3u) == 0) {
    /* rows are word aligned: each lane decodes four bases and stores one 32-bit word */
    for (uint32_t j = 4 * lane; j < max_len; j += 128) {
      uint32_t word = 0;
#pragma unroll
      for (uint32_t t = 0; t < 4; ++t) {
        if (j + t < m) {
          const uint64_t p = base + j + t;
          word |= ((0x54474341u >> (8 * ((packed[p >> 4] >> (2 * (p & 15))) & 3u))) & 0xFFu) << (8 * t);
        }
      }
      *reinterpret_cast<uint32_t*>(dst + j) = word;
    }
  } else {
    for (uint32_t j = lane; j < max_len; j += 32) {
      uint8_t ch = 0;
      if (j < m) {
        const uint64_t p = base + j;
        ch = (uint8_t)((0x54474341u >> (8 * ((packed[p >> 4] >> (2 * (p & 15))) & 3u))) & 0xFFu); /* "ACGT" */
      }
      dst[j] = ch;
    }
  }
  if (lane == 0) out_len[i] = r < num_strings ? len : 0xFFFFFFFFu;
}

int rs_index_set_reads(rs_index* ix, const char* text, const uint64_t* offsets, uint64_t num_strings) {
  if (!ix || !offsets || (!text && num_strings && offsets[num_strings])) return fail(RS_STATUS_ARG, "null argument");
  if (num_strings != ix->dev.num_strings)
    return fail(RS_STATUS_ARG, "read table has %llu reads, the index holds %llu", (unsigned long long)num_strings,
                (unsigned long long)ix->dev.num_strings);
  for (uint64_t r = 0; r < num_strings; ++r)
    if (offsets[r + 1] < offsets[r]) return fail(RS_STATUS_ARG, "read offsets are not ascending at read %llu", (unsigned long long)r);
  const uint64_t nbases = offsets[num_strings];
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  if (ix->d_reads) {
    cudaFree(ix->d_reads);
    cudaFree(ix->d_read_off);
    ix->hbm_bytes -= ix->reads_bytes;
    ix->d_reads = nullptr;
    ix->d_read_off = nullptr;
    ix->reads_bytes = 0;
  }
  const uint64_t nwords = (nbases + 15) / 16;
  uint8_t* d_text = nullptr;
  unsigned int* d_bad = nullptr;
  CU(cudaMalloc((void**)&ix->d_reads, (nwords ? nwords : 1) * 4));
  CU(cudaMalloc((void**)&ix->d_read_off, (num_strings + 1) * 8));
  CU(cudaMalloc((void**)&d_text, nbases ? nbases : 1));
  CU(cudaMalloc((void**)&d_bad, 4));
  cudaStream_t st = ix->stream;
  CU(cudaMemsetAsync(d_bad, 0, 4, st));
  CU(cudaMemcpyAsync(d_text, text, nbases, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ix->d_read_off, offsets, (num_strings + 1) * 8, cudaMemcpyHostToDevice, st));
  if (nwords) k_pack_reads<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(d_text, nbases, ix->d_reads, d_bad);
  unsigned int bad = 0;
  CU(cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  cudaFree(d_text);
  cudaFree(d_bad);
  if (bad) {
    cudaFree(ix->d_reads);
    cudaFree(ix->d_read_off);
    ix->d_reads = nullptr;
    ix->d_read_off = nullptr;
    return fail(RS_STATUS_ARG, "the read table holds a byte outside ACGT (such reads are not in the BWT)");
  }
  ix->reads_bytes = nwords * 4 + (num_strings + 1) * 8;
  ix->hbm_bytes += ix->reads_bytes;
  return RS_STATUS_OK;
}

int rs_index_load_reads_file(rs_index* ix, const char* path) {
  if (!ix || !path) return fail(RS_STATUS_ARG, "null argument");
  std::string text;
  std::vector<uint64_t> offsets;
  const int rc = rs_parse_etl_reads(path, text, offsets);
  if (rc == RS_ERR_OPEN) return fail(RS_STATUS_OPEN, "cannot open %s", path);
  if (rc != RS_OK) return fail(RS_STATUS_ARG, "%s is not a `read / codes` file", path);
  return rs_index_set_reads(ix, text.data(), offsets.data(), offsets.size() - 1);
}

int rs_reads_batch_dev(const rs_index* ix, const void* d_read_index, size_t n, uint32_t max_len, void* d_out,
                       void* d_out_len, void* stream) {
  if (!ix || (n && (!d_read_index || !d_out || !d_out_len))) return fail(RS_STATUS_ARG, "null argument");
  if (!ix->d_reads) return fail(RS_STATUS_NO_TABLE, "no read table attached");
  if (max_len < 1) return fail(RS_STATUS_ARG, "max_len must be >= 1");
  if (!n) return RS_STATUS_OK;
  k_unpack_reads<<<(unsigned)((n * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      ix->d_reads, ix->d_read_off, ix->dev.num_strings, (const uint64_t*)d_read_index, n, max_len, (uint8_t*)d_out,
      (uint32_t*)d_out_len);
  CU(cudaGetLastError());
  return RS_STATUS_OK;
}

int rs_reads_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint32_t max_len, char* out, uint32_t* out_len) {
  if (!ix || (n && (!read_index || !out || !out_len))) return fail(RS_STATUS_ARG, "null argument");
  if (!ix->d_reads) return fail(RS_STATUS_NO_TABLE, "no read table attached");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  CU(ix->s_in.need(n * 8));
  CU(ix->s_out.need(n * (size_t)max_len));
  CU(ix->s_out2.need(n * 4));
  CU(cudaMemcpyAsync(ix->s_in.p, read_index, n * 8, cudaMemcpyHostToDevice, st));
  const int rc = rs_reads_batch_dev(ix, ix->s_in.p, n, max_len, ix->s_out.p, ix->s_out2.p, st);
  if (rc != RS_STATUS_OK) return rc;
  CU(cudaMemcpyAsync(out, ix->s_out.p, n * (size_t)max_len, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(out_len, ix->s_out2.p, n * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}
