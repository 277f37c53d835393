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
  std::vector<uint8_t> codes;
  const int rc = rs_parse_etl_samples(path, size_of_sample, has_other_meta_data, offsets, codes);
  if (rc == RS_ERR_OPEN) return fail(RS_STATUS_OPEN, "cannot open %s", path);
  if (rc != RS_OK) return fail(RS_STATUS_ARG, "%s is not a `read / codes` file for size_of_sample %d", path, size_of_sample);
  return rs_index_set_samples(ix, offsets.data(), codes.data(), offsets.size() - 1, size_of_sample);
}

int rs_index_set_samples(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, uint64_t num_strings,
                         int size_of_sample) {
  if (!ix || !offsets) return fail(RS_STATUS_ARG, "null argument");
  if (size_of_sample < 1 || size_of_sample > 4) return fail(RS_STATUS_ARG, "size_of_sample must be 1..4");
  if (num_strings != ix->dev.num_strings)
    return fail(RS_STATUS_ARG, "sample table has %llu reads, the index holds %llu", (unsigned long long)num_strings,
                (unsigned long long)ix->dev.num_strings);
  const uint64_t ncodes = offsets[num_strings];
  if (ncodes && !codes) return fail(RS_STATUS_ARG, "null codes");
  for (uint64_t r = 0; r < num_strings; ++r)
    if (offsets[r + 1] < offsets[r]) return fail(RS_STATUS_ARG, "sample offsets are not ascending at read %llu", (unsigned long long)r);
  /* largest sample number in the table (create_samples_hash.cpp:41-50: base-223 digits + 33, 1-based) */
  uint64_t max_sample = 0;
  for (uint64_t t = 0; t < ncodes; ++t) {
    uint64_t v = 0;
    for (int i = 0; i < size_of_sample; ++i) {
      const uint8_t ch = codes[t * size_of_sample + i];
      if (ch < 33) return fail(RS_STATUS_ARG, "byte %u in sample code %llu is not a base-223 digit", ch, (unsigned long long)t);
      v = v * 223 + (uint64_t)(ch - 33);
    }
    if (v == 0) return fail(RS_STATUS_ARG, "sample code %llu decodes to 0 (codes are 1-based)", (unsigned long long)t);
    if (v > max_sample) max_sample = v;
  }
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  if (ix->d_samp_off) {
    cudaFree(ix->d_samp_off);
    cudaFree(ix->d_samp_codes);
    ix->d_samp_off = nullptr;
    ix->d_samp_codes = nullptr;
  }
  const size_t ob = (num_strings + 1) * 8, cb = ncodes * (size_t)size_of_sample;
  CU(cudaMalloc((void**)&ix->d_samp_off, ob));
  CU(cudaMalloc((void**)&ix->d_samp_codes, cb ? cb : 8));
  CU(cudaMemcpy(ix->d_samp_off, offsets, ob, cudaMemcpyHostToDevice));
  if (cb) CU(cudaMemcpy(ix->d_samp_codes, codes, cb, cudaMemcpyHostToDevice));
  ix->size_of_sample = size_of_sample;
  ix->num_samples = (uint32_t)max_sample;
  ix->hbm_bytes += ob + cb;
  return RS_STATUS_OK;
}

int rs_samples_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint64_t* out_offsets,
                     uint64_t capacity, uint8_t* out_codes) {
  if (!ix || !out_offsets || (n && !read_index)) return fail(RS_STATUS_ARG, "null argument");
  if (!ix->d_samp_off) return fail(RS_STATUS_NO_TABLE, "no sample table attached");
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
  CU(tmp.get((void**)&d_codes, total * sos));
  CU(rs_launch_sample_gather(ix->d_samp_off, ix->d_samp_codes, sos, ix->dev.num_strings, d_ri, n, d_off, d_codes, st));
  CU(cudaMemcpyAsync(out_codes, d_codes, total * sos, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
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
