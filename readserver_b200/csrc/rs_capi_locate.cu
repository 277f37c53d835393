/* C ABI, second part: count+locate pipeline, sample table, carrier sets (include/readserver_b200.h). */
#include "rs_internal.h"

/* intervals -> hit offsets (device); returns total hits */
static int hits_from_intervals(const uint64_t* d_lower, const uint64_t* d_upper, size_t Q, uint64_t max_hits,
                               uint64_t* d_hit_offsets, uint64_t* total, Tmp& tmp) {
  uint64_t* d_counts;
  void* d_scan;
  size_t scan_bytes = 0;
  CU(tmp.get((void**)&d_counts, (Q + 1) * 8));
  CU(rs_scan_bytes(Q, &scan_bytes));
  CU(tmp.get(&d_scan, scan_bytes));
  CU(rs_launch_hit_counts(d_lower, d_upper, Q, max_hits, d_counts, tmp.st));
  CU(rs_exclusive_scan(d_counts, d_hit_offsets, Q, d_scan, scan_bytes, tmp.st));
  CU(cudaMemcpyAsync(total, d_hit_offsets + Q, 8, cudaMemcpyDeviceToHost, tmp.st));
  CU(cudaStreamSynchronize(tmp.st));
  return RS_STATUS_OK;
}

int rs_locate_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, uint64_t max_hits,
                        uint32_t max_steps, void* d_hit_offsets, uint64_t capacity, void* d_rows, void* d_lexrank,
                        void* d_read_index, void* d_steps, uint64_t* total_hits, void* stream) {
  if (!ix || !d_hit_offsets || !total_hits || (Q && !d_kmers)) return fail(RS_STATUS_ARG, "null argument");
  if (k < 1) return fail(RS_STATUS_ARG, "k must be >= 1 (empty query: SURVEY.md Q3)");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  if (d_read_index && !ix->d_sai) return fail(RS_STATUS_NO_TABLE, "read_index requested but no .sai table attached");
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = (cudaStream_t)stream;
  Tmp tmp(st);
  uint64_t *d_lo, *d_up;
  CU(tmp.get((void**)&d_lo, Q * 8));
  CU(tmp.get((void**)&d_up, Q * 8));
  CU(rs_launch_find_interval(ix->dev, (const uint8_t*)d_kmers, nullptr, (uint32_t)k, Q, d_lo, d_up, 0, 0,
                             (unsigned long long*)ix->d_stats, st));
  const int rc = hits_from_intervals(d_lo, d_up, Q, max_hits, (uint64_t*)d_hit_offsets, total_hits, tmp);
  if (rc != RS_STATUS_OK) return rc;
  const uint64_t H = *total_hits;
  if (H > capacity)
    return fail(RS_STATUS_CAPACITY, "%llu hits do not fit the %llu provided", (unsigned long long)H,
                (unsigned long long)capacity);
  if (!H || (!d_rows && !d_lexrank && !d_read_index && !d_steps)) return RS_STATUS_OK;
  uint64_t* rows = (uint64_t*)d_rows;
  if (!rows) CU(tmp.get((void**)&rows, H * 8));
  CU(rs_launch_expand_rows(d_lo, (const uint64_t*)d_hit_offsets, Q, H, rows, nullptr, st));
  if (d_lexrank || d_read_index || d_steps)
    CU(rs_launch_extract_prefix(ix->dev, rows, H, max_steps, (uint64_t*)d_lexrank, (uint32_t*)d_steps, ix->d_sai,
                                (uint64_t*)d_read_index, nullptr, (unsigned long long*)ix->d_stats, st));
  return RS_STATUS_OK;
}

int rs_locate_batch(const rs_index* ix, const char* kmers, int k, size_t Q, uint64_t max_hits, uint32_t max_steps,
                    uint64_t* hit_offsets, uint64_t capacity, uint64_t* rows, uint64_t* lexrank, uint64_t* read_index,
                    uint32_t* steps) {
  if (!ix || !hit_offsets || (Q && !kmers)) return fail(RS_STATUS_ARG, "null argument");
  if (k < 1) return fail(RS_STATUS_ARG, "k must be >= 1 (empty query: SURVEY.md Q3)");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  Tmp tmp(st);
  uint8_t* d_k;
  uint64_t *d_off, *d_rows = nullptr, *d_lex = nullptr, *d_ri = nullptr;
  uint32_t* d_st = nullptr;
  CU(tmp.get((void**)&d_k, Q * (size_t)k));
  CU(tmp.get((void**)&d_off, (Q + 1) * 8));
  if (capacity) {
    if (rows) CU(tmp.get((void**)&d_rows, capacity * 8));
    if (lexrank) CU(tmp.get((void**)&d_lex, capacity * 8));
    if (read_index) CU(tmp.get((void**)&d_ri, capacity * 8));
    if (steps) CU(tmp.get((void**)&d_st, capacity * 4));
  }
  CU(cudaMemcpyAsync(d_k, kmers, Q * (size_t)k, cudaMemcpyHostToDevice, st));
  uint64_t H = 0;
  const int rc = rs_locate_batch_dev(ix, d_k, k, Q, max_hits, max_steps, d_off, capacity, d_rows, d_lex, d_ri, d_st, &H, st);
  if (rc != RS_STATUS_OK && rc != RS_STATUS_CAPACITY) return rc;
  CU(cudaMemcpyAsync(hit_offsets, d_off, (Q + 1) * 8, cudaMemcpyDeviceToHost, st));
  if (rc == RS_STATUS_OK && H) {
    if (rows) CU(cudaMemcpyAsync(rows, d_rows, H * 8, cudaMemcpyDeviceToHost, st));
    if (lexrank) CU(cudaMemcpyAsync(lexrank, d_lex, H * 8, cudaMemcpyDeviceToHost, st));
    if (read_index) CU(cudaMemcpyAsync(read_index, d_ri, H * 8, cudaMemcpyDeviceToHost, st));
    if (steps) CU(cudaMemcpyAsync(steps, d_st, H * 4, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(st));
  return rc;
}

/* ---- sample table ---- */
int rs_index_load_samples_file(rs_index* ix, const char* path, int size_of_sample, int has_other_meta_data) {
  if (!ix || !path) return fail(RS_STATUS_ARG, "null argument");
  std::vector<uint64_t> offsets;
  std::vector<uint8_t> codes, meta;
  int rc;
  try {
    rc = rs_parse_etl_samples(path, size_of_sample, has_other_meta_data, offsets, codes, has_other_meta_data ? &meta : nullptr);
  } catch (const std::exception&) {
    return fail(RS_STATUS_ARG, "out of host memory while reading %s", path);
  }
  if (rc == RS_ERR_OPEN) return fail(RS_STATUS_OPEN, "cannot open %s", path);
  if (rc != RS_OK) return fail(RS_STATUS_ARG, "%s is not a `read / codes` file for size_of_sample %d", path, size_of_sample);
  return rs_index_set_samples_meta(ix, offsets.data(), codes.data(), has_other_meta_data ? meta.data() : nullptr,
                                   offsets.size() - 1, size_of_sample);
}

/* validation of a sample table on the device: offsets ascending, every code byte a base-223 digit (>= 33),
 * no code decoding to 0; max[0] = largest sample number (create_samples_hash.cpp:41-50: 1-based) */
__global__ void k_check_offsets(const uint64_t* __restrict__ off, uint64_t n, unsigned int* __restrict__ bad) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride)
    if (off[r + 1] < off[r]) atomicOr(bad, 1u);
}
__global__ void k_check_codes(const uint8_t* __restrict__ codes, uint64_t ncodes, uint32_t sos, uint32_t stride_bytes,
                              unsigned int* __restrict__ bad, unsigned long long* __restrict__ maxv) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long m = 0;
  unsigned int b = 0;
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < ncodes; t += stride) {
    unsigned long long v = 0;
    for (uint32_t i = 0; i < sos; ++i) {
      const uint32_t ch = codes[t * stride_bytes + i];
      if (ch < 33) b |= 2u;
      v = v * 223ull + (unsigned long long)(ch - 33u);
    }
    if (v == 0) b |= 4u;
    if (v > m) m = v;
  }
  if (b) atomicOr(bad, b);
  if (m) atomicMax(maxv, m);
}

static int adopt_samples(rs_index* ix, uint64_t* d_off, uint8_t* d_codes, uint8_t* d_meta, uint64_t num_strings, int sos,
                         uint64_t ncodes) {
  /* takes ownership of the device arrays (frees them on failure) */
  unsigned int* d_bad = nullptr;
  unsigned long long* d_max = nullptr;
  unsigned int bad = 0;
  unsigned long long maxv = 0;
  cudaError_t e = cudaMalloc((void**)&d_bad, 4);
  if (e == cudaSuccess) e = cudaMalloc((void**)&d_max, 8);
  if (e == cudaSuccess) e = cudaMemset(d_bad, 0, 4);
  if (e == cudaSuccess) e = cudaMemset(d_max, 0, 8);
  if (e == cudaSuccess) {
    if (num_strings) k_check_offsets<<<148 * 4, 256>>>(d_off, num_strings, d_bad);
    if (ncodes) k_check_codes<<<148 * 8, 256>>>(d_codes, ncodes, (uint32_t)sos, (uint32_t)sos, d_bad, d_max);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(&bad, d_bad, 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(&maxv, d_max, 8, cudaMemcpyDeviceToHost);
  if (d_bad) cudaFree(d_bad);
  if (d_max) cudaFree(d_max);
  if (e != cudaSuccess || bad) {
    cudaFree(d_off);
    cudaFree(d_codes);
    if (d_meta) cudaFree(d_meta);
    if (e != cudaSuccess) return fail(RS_STATUS_CUDA, "sample table: %s", cudaGetErrorString(e));
    if (bad & 1u) return fail(RS_STATUS_ARG, "sample offsets are not ascending");
    if (bad & 2u) return fail(RS_STATUS_ARG, "a sample code holds a byte that is not a base-223 digit (< 33)");
    return fail(RS_STATUS_ARG, "a sample code decodes to 0 (codes are 1-based)");
  }
  if (ix->d_samp_off) {
    cudaFree(ix->d_samp_off);
    cudaFree(ix->d_samp_codes);
    if (ix->d_samp_meta) cudaFree(ix->d_samp_meta);
    ix->hbm_bytes -= ix->samp_bytes;
  }
  ix->d_samp_off = d_off;
  ix->d_samp_codes = d_codes;
  ix->d_samp_meta = d_meta;
  ix->size_of_sample = sos;
  ix->num_samples = (uint32_t)maxv;
  ix->samp_bytes = (num_strings + 1) * 8 + ncodes * (uint64_t)sos + (d_meta ? ncodes * 2 : 0);
  ix->hbm_bytes += ix->samp_bytes;
  return RS_STATUS_OK;
}

static int set_samples_any(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, const uint8_t* meta,
                           uint64_t num_strings, int size_of_sample, uint64_t ncodes, cudaMemcpyKind kind) {
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  uint64_t* d_off = nullptr;
  uint8_t *d_codes = nullptr, *d_meta = nullptr;
  const size_t ob = (num_strings + 1) * 8, cb = ncodes * (size_t)size_of_sample;
  cudaError_t e = cudaMalloc((void**)&d_off, ob);
  if (e == cudaSuccess) e = cudaMalloc((void**)&d_codes, cb ? cb : 8);
  if (e == cudaSuccess && meta) e = cudaMalloc((void**)&d_meta, ncodes ? ncodes * 2 : 8);
  if (e == cudaSuccess) e = cudaMemcpy(d_off, offsets, ob, kind);
  if (e == cudaSuccess && cb) e = cudaMemcpy(d_codes, codes, cb, kind);
  if (e == cudaSuccess && meta && ncodes) e = cudaMemcpy(d_meta, meta, ncodes * 2, kind);
  if (e != cudaSuccess) {
    if (d_off) cudaFree(d_off);
    if (d_codes) cudaFree(d_codes);
    if (d_meta) cudaFree(d_meta);
    return fail(RS_STATUS_CUDA, "sample table: %s", cudaGetErrorString(e));
  }
  return adopt_samples(ix, d_off, d_codes, d_meta, num_strings, size_of_sample, ncodes);
}

static int samples_args(rs_index* ix, const void* offsets, uint64_t num_strings, int size_of_sample) {
  if (!ix || !offsets) return fail(RS_STATUS_ARG, "null argument");
  if (size_of_sample < 1 || size_of_sample > 4) return fail(RS_STATUS_ARG, "size_of_sample must be 1..4");
  if (num_strings != ix->dev.num_strings)
    return fail(RS_STATUS_ARG, "sample table has %llu reads, the index holds %llu", (unsigned long long)num_strings,
                (unsigned long long)ix->dev.num_strings);
  return RS_STATUS_OK;
}

int rs_index_set_samples_meta(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, const uint8_t* meta,
                              uint64_t num_strings, int size_of_sample) {
  const int rc = samples_args(ix, offsets, num_strings, size_of_sample);
  if (rc != RS_STATUS_OK) return rc;
  const uint64_t ncodes = offsets[num_strings];
  if (ncodes && !codes) return fail(RS_STATUS_ARG, "null codes");
  return set_samples_any(ix, offsets, codes, meta, num_strings, size_of_sample, ncodes, cudaMemcpyHostToDevice);
}

int rs_index_set_samples(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, uint64_t num_strings,
                         int size_of_sample) {
  return rs_index_set_samples_meta(ix, offsets, codes, nullptr, num_strings, size_of_sample);
}

int rs_index_set_samples_dev(rs_index* ix, const void* d_offsets, const void* d_codes, uint64_t num_strings,
                             int size_of_sample) {
  const int rc = samples_args(ix, d_offsets, num_strings, size_of_sample);
  if (rc != RS_STATUS_OK) return rc;
  CU(cudaSetDevice(ix->device));
  uint64_t ncodes = 0;
  CU(cudaMemcpy(&ncodes, (const uint64_t*)d_offsets + num_strings, 8, cudaMemcpyDeviceToHost));
  if (ncodes && !d_codes) return fail(RS_STATUS_ARG, "null codes");
  return set_samples_any(ix, (const uint64_t*)d_offsets, (const uint8_t*)d_codes, nullptr, num_strings, size_of_sample,
                         ncodes, cudaMemcpyDeviceToDevice);
}

int rs_samples_meta_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint64_t* out_offsets,
                          uint64_t capacity, uint8_t* out_codes, uint8_t* out_meta) {
  if (!ix || !out_offsets || (n && !read_index)) return fail(RS_STATUS_ARG, "null argument");
  if (!ix->d_samp_off) return fail(RS_STATUS_NO_TABLE, "no sample table attached");
  if (out_meta && !ix->d_samp_meta) return fail(RS_STATUS_NO_TABLE, "the sample table holds no count / lane metadata");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  Tmp tmp(st);
  const uint32_t sos = (uint32_t)ix->size_of_sample;
  uint64_t *d_ri, *d_cnt, *d_off;
  void* d_scan;
  size_t scan_bytes = 0;
  CU(tmp.get((void**)&d_ri, n * 8));
  CU(tmp.get((void**)&d_cnt, (n + 1) * 8));
  CU(tmp.get((void**)&d_off, (n + 1) * 8));
  CU(rs_scan_bytes(n, &scan_bytes));
  CU(tmp.get(&d_scan, scan_bytes));
  CU(cudaMemcpyAsync(d_ri, read_index, n * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_sample_counts(ix->d_samp_off, ix->dev.num_strings, d_ri, n, d_cnt, st));
  CU(rs_exclusive_scan(d_cnt, d_off, n, d_scan, scan_bytes, st));
  CU(cudaMemcpyAsync(out_offsets, d_off, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const uint64_t total = out_offsets[n];
  if (total > capacity || (total && !out_codes))
    return fail(RS_STATUS_CAPACITY, "%llu sample codes do not fit the %llu provided", (unsigned long long)total,
                (unsigned long long)capacity);
  if (!total) return RS_STATUS_OK;
  uint8_t* d_codes;
  CU(tmp.get((void**)&d_codes, total * (sos > 2 ? sos : 2)));
  CU(rs_launch_sample_gather(ix->d_samp_off, ix->d_samp_codes, sos, ix->dev.num_strings, d_ri, n, d_off, d_codes, st));
  CU(cudaMemcpyAsync(out_codes, d_codes, total * sos, cudaMemcpyDeviceToHost, st));
  if (out_meta) { /* the count and lane bytes travel beside the codes: same offsets, 2 bytes per code */
    CU(rs_launch_sample_gather(ix->d_samp_off, ix->d_samp_meta, 2, ix->dev.num_strings, d_ri, n, d_off, d_codes, st));
    CU(cudaMemcpyAsync(out_meta, d_codes, total * 2, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

int rs_samples_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint64_t* out_offsets,
                     uint64_t capacity, uint8_t* out_codes) {
  return rs_samples_meta_batch(ix, read_index, n, out_offsets, capacity, out_codes, nullptr);
}

int rs_carriers_batch(const rs_index* ix, const uint64_t* hit_offsets, const uint64_t* read_index, size_t Q,
                      uint64_t* out_offsets, uint64_t capacity, uint8_t* out_codes) {
  if (!ix || !hit_offsets || !out_offsets) return fail(RS_STATUS_ARG, "null argument");
  if (!ix->d_samp_off) return fail(RS_STATUS_NO_TABLE, "no sample table attached");
  const uint64_t H = hit_offsets[Q];
  if (H && !read_index) return fail(RS_STATUS_ARG, "null read_index");
  if ((size_t)8 * ((ix->num_samples + 31) / 32) * 4 > 48 * 1024)
    return fail(RS_STATUS_TOO_LARGE, "carrier sets support at most 49152 samples (table has %u)", ix->num_samples);
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  cudaStream_t st = ix->stream;
  Tmp tmp(st);
  const uint32_t sos = (uint32_t)ix->size_of_sample;
  uint64_t *d_hoff, *d_ri, *d_cnt, *d_off;
  void* d_scan;
  size_t scan_bytes = 0;
  CU(tmp.get((void**)&d_hoff, (Q + 1) * 8));
  CU(tmp.get((void**)&d_ri, H * 8));
  CU(tmp.get((void**)&d_cnt, (Q + 1) * 8));
  CU(tmp.get((void**)&d_off, (Q + 1) * 8));
  CU(rs_scan_bytes(Q, &scan_bytes));
  CU(tmp.get(&d_scan, scan_bytes));
  CU(cudaMemcpyAsync(d_hoff, hit_offsets, (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  if (H) CU(cudaMemcpyAsync(d_ri, read_index, H * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_carriers(ix->d_samp_off, ix->d_samp_codes, sos, ix->dev.num_strings, ix->num_samples, d_hoff, d_ri, Q, 0,
                        d_cnt, nullptr, nullptr, st));
  CU(rs_exclusive_scan(d_cnt, d_off, Q, d_scan, scan_bytes, st));
  CU(cudaMemcpyAsync(out_offsets, d_off, (Q + 1) * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const uint64_t total = out_offsets[Q];
  if (total > capacity || (total && !out_codes))
    return fail(RS_STATUS_CAPACITY, "%llu carrier codes do not fit the %llu provided", (unsigned long long)total,
                (unsigned long long)capacity);
  if (!total) return RS_STATUS_OK;
  uint8_t* d_codes;
  CU(tmp.get((void**)&d_codes, total * sos));
  CU(rs_launch_carriers(ix->d_samp_off, ix->d_samp_codes, sos, ix->dev.num_strings, ix->num_samples, d_hoff, d_ri, Q, 1,
                        nullptr, d_off, d_codes, st));
  CU(cudaMemcpyAsync(out_codes, d_codes, total * sos, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}
