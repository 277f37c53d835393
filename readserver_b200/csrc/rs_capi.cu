/* C ABI of the engine (include/readserver_b200.h): index lifecycle, host<->device staging and
 * kernel launches.  No torch types, no CPU compute path. */
#include "rs_internal.h"

#include <stdlib.h>

#include <new>

#include "rs_encode.h"

static thread_local std::string g_err;

int rs_fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

const char* rs_last_error(void) { return g_err.c_str(); }


static void set_tail(rs_index* ix, bool on) { rs_view_set_tail(ix->init_occ_ref, on, &ix->dev); }

/* (Re)build the k-mer suffix table with t characters (t < 2: drop it).  Must run after any change of
 * the tail-emulation switch: the table holds reference-semantics or exact intervals accordingly. */
static int build_ftab(rs_index* ix, uint32_t t) {
  if (ix->d_ftab) {
    CU(cudaStreamSynchronize(ix->stream));
    cudaFree(ix->d_ftab);
    ix->hbm_bytes -= (16ull << (2 * ix->dev.ftab_k));
    ix->d_ftab = nullptr;
  }
  ix->dev.ftab = nullptr;
  ix->dev.ftab_k = 0;
  if (t < 2) return RS_STATUS_OK;
  if (t > 12) return fail(RS_STATUS_ARG, "prefix table length must be <= 12");
  const size_t bytes = 16ull << (2 * t);
  CU(cudaMalloc((void**)&ix->d_ftab, bytes));
  CU(rs_build_ftab(ix->dev, t, ix->d_ftab, ix->stream));
  CU(cudaStreamSynchronize(ix->stream));
  ix->dev.ftab = ix->d_ftab;
  ix->dev.ftab_k = t;
  ix->hbm_bytes += bytes;
  return RS_STATUS_OK;
}

/* Dense rank blocks (rs_layout.h): 0.5 bytes per symbol beside the run blocks; rank and access then need one memory
 * access and a few instructions.  RS_B200_DENSE = 0 never, 1 always; default: when they cost at most twice the run
 * blocks (runs of ~7 symbols or shorter) and every base occurs fewer than 2^32 times. */
static int drop_dense(rs_index* ix) {
  if (ix->d_dense) {
    CU(cudaStreamSynchronize(ix->stream));
    cudaFree(ix->d_dense);
    ix->hbm_bytes -= ix->dense_bytes;
    ix->d_dense = nullptr;
    ix->dense_bytes = 0;
  }
  ix->dev.dense = nullptr;
  return RS_STATUS_OK;
}
static int build_dense(rs_index* ix) {
  int rc = drop_dense(ix);
  if (rc != RS_STATUS_OK) return rc;
  for (int c = 1; c < RS_NSYM; ++c)
    if (ix->dev.tot[c] >> 32) return fail(RS_STATUS_TOO_LARGE, "dense rank blocks need fewer than 2^32 occurrences of each base");
  const uint64_t nd = (ix->dev.n + RS_DENSE_SYMS - 1) >> RS_DENSE_SHIFT;
  unsigned int* d_bad = nullptr;
  unsigned int bad = 0;
  CU(cudaMalloc(&ix->d_dense, (nd ? nd : 1) * 64));
  cudaError_t e = cudaMalloc((void**)&d_bad, 4);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_bad, 0, 4, ix->stream);
  if (e == cudaSuccess) e = rs_build_dense(ix->dev, nd, ix->d_dense, d_bad, ix->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, ix->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ix->stream);
  if (d_bad) cudaFree(d_bad);
  if (e != cudaSuccess || bad) {
    cudaFree(ix->d_dense);
    ix->d_dense = nullptr;
    if (e != cudaSuccess) return fail(RS_STATUS_CUDA, "dense rank blocks: %s", cudaGetErrorString(e));
    return fail(RS_STATUS_TOO_LARGE, "dense rank blocks: a count does not fit 32 bits");
  }
  ix->dense_bytes = nd * 64;
  ix->hbm_bytes += ix->dense_bytes;
  ix->dev.dense = ix->d_dense;
  return RS_STATUS_OK;
}
static bool auto_dense(const rs_index* ix) {
  const char* e = getenv("RS_B200_DENSE");
  if (e && e[0] == '0') return false;
  for (int c = 1; c < RS_NSYM; ++c)
    if (ix->dev.tot[c] >> 32) return false;
  if (e && e[0] == '1') return true;
  return ix->dev.n / 2 <= 2 * ix->dev.nblocks * (uint64_t)RS_BLOCK_BYTES;
}

int rs_index_set_dense(rs_index* ix, int on) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  return on ? build_dense(ix) : drop_dense(ix);
}

/* (Re)build or drop the locate table.  Like the suffix table it holds the answers of the semantics in force
 * (reference tail behaviour or exact), so the tail switch rebuilds it. */
static int drop_locate(rs_index* ix) {
  if (ix->d_loc) {
    CU(cudaStreamSynchronize(ix->stream));
    cudaFree(ix->d_loc);
    cudaFree(ix->d_loc_fin);
    if (ix->d_loc_mark) cudaFree(ix->d_loc_mark);
    if (ix->d_loc_rank) cudaFree(ix->d_loc_rank);
    ix->hbm_bytes -= ix->loc_bytes;
    ix->d_loc = ix->d_loc_fin = ix->d_loc_mark = ix->d_loc_rank = nullptr;
    ix->loc_bytes = 0;
  }
  ix->dev.loc = ix->dev.loc_fin = ix->dev.loc_mark = ix->dev.loc_rank = nullptr;
  ix->dev.loc_every = 0;
  return RS_STATUS_OK;
}
static int build_locate(rs_index* ix, uint32_t max_len, uint32_t every) {
  const int rc = drop_locate(ix);
  if (rc != RS_STATUS_OK) return rc;
  if (ix->dev.num_strings >> 40) return fail(RS_STATUS_TOO_LARGE, "locate table: more than 2^40 reads");
  const uint64_t n = ix->dev.n, ns = ix->dev.num_strings;
  cudaStream_t st = ix->stream;
  uint64_t bytes = 0;
  auto alloc = [&](uint64_t** p, uint64_t count) {
    const cudaError_t e = cudaMalloc((void**)p, (count ? count : 1) * 8);
    if (e == cudaSuccess) bytes += (count ? count : 1) * 8;
    return e;
  };
  auto bail = [&](const char* what) {
    if (ix->d_loc) cudaFree(ix->d_loc);
    if (ix->d_loc_fin) cudaFree(ix->d_loc_fin);
    if (ix->d_loc_mark) cudaFree(ix->d_loc_mark);
    if (ix->d_loc_rank) cudaFree(ix->d_loc_rank);
    ix->d_loc = ix->d_loc_fin = ix->d_loc_mark = ix->d_loc_rank = nullptr;
    cudaGetLastError();
    return fail(RS_STATUS_CUDA, "locate table: %s", what);
  };
  if (alloc(&ix->d_loc_fin, ns) != cudaSuccess) return bail("out of device memory");
  if (every <= 1) {
    if (alloc(&ix->d_loc, n) != cudaSuccess) return bail("out of device memory");
    if (rs_build_locate_table(ix->dev, max_len, ix->d_loc, ix->d_loc_fin, st) != cudaSuccess) return bail("build failed");
  } else {
    const uint64_t nwords = (n + 63) / 64;
    if (alloc(&ix->d_loc_mark, 2 * nwords) != cudaSuccess) return bail("out of device memory");
    uint64_t *d_counts = nullptr, *d_before = nullptr;
    void* d_scan = nullptr;
    size_t scan_bytes = 0;
    if (rs_scan_bytes(nwords, &scan_bytes) != cudaSuccess || cudaMalloc((void**)&d_counts, (nwords + 1) * 8) != cudaSuccess ||
        cudaMalloc((void**)&d_before, (nwords + 1) * 8) != cudaSuccess || cudaMalloc(&d_scan, scan_bytes ? scan_bytes : 8) != cudaSuccess) {
      if (d_counts) cudaFree(d_counts);
      if (d_before) cudaFree(d_before);
      return bail("out of device memory");
    }
    uint64_t nsamples = 0;
    cudaError_t e = rs_locate_mark(ix->dev, max_len, every, ix->d_loc_mark, ix->d_loc_fin, st);
    if (e == cudaSuccess) e = rs_locate_mark_counts(ix->d_loc_mark, n, d_counts, st);
    if (e == cudaSuccess) e = rs_exclusive_scan(d_counts, d_before, nwords, d_scan, scan_bytes, st);
    if (e == cudaSuccess) e = rs_locate_mark_ranks(ix->d_loc_mark, n, d_before, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nsamples, d_before + nwords, 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_counts);
    cudaFree(d_before);
    cudaFree(d_scan);
    if (e != cudaSuccess) return bail(cudaGetErrorString(e));
    if (alloc(&ix->d_loc, nsamples) != cudaSuccess) return bail("out of device memory");
    if (rs_locate_fill(ix->dev, max_len, every, ix->d_loc_mark, ix->d_loc, nsamples, st) != cudaSuccess)
      return bail("build failed");
  }
  if (cudaStreamSynchronize(st) != cudaSuccess) return bail("build failed");
  ix->dev.loc = ix->d_loc;
  ix->dev.loc_fin = ix->d_loc_fin;
  ix->dev.loc_mark = ix->d_loc_mark;
  ix->dev.loc_rank = ix->d_loc_rank;
  ix->dev.loc_every = every <= 1 ? 1 : every;
  ix->loc_max_len = max_len;
  ix->loc_every = every;
  ix->loc_bytes = bytes;
  ix->hbm_bytes += bytes;
  return RS_STATUS_OK;
}

int rs_index_build_locate_table(rs_index* ix, uint32_t max_read_len) { return rs_index_build_locate_samples(ix, max_read_len, 1); }

int rs_index_build_locate_samples(rs_index* ix, uint32_t max_read_len, uint32_t every) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  if (max_read_len < 1 || max_read_len >= (1u << 24) - 1) return fail(RS_STATUS_ARG, "max_read_len must be in [1, 2^24 - 2]");
  if (every < 1) return fail(RS_STATUS_ARG, "every must be >= 1");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  return build_locate(ix, max_read_len, every);
}

int rs_index_drop_locate_table(rs_index* ix) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  return drop_locate(ix);
}

static uint32_t auto_ftab_k(const rs_index* ix) {
  const char* e = getenv("RS_B200_FTAB");
  if (e) return (uint32_t)atoi(e);
  /* largest t <= 12 whose table is no larger than the run blocks -- four times the run blocks when the index has
   * dense rank blocks (short runs: the run blocks are then a poor measure of the index).  One 16-byte read per query
   * replaces t - 1 steps of two block reads each, so the table pays even when it is larger than L2.  configs[1]
   * (run blocks 111 MB, M q/s): t = 8: 1540, 9: 1598, 10: 1674, 11: 1802, 12: 1842 in round 1, 11: 1428 / 12: 1258 on
   * this round's box; configs[2] (64 shards of 73 MB run blocks + 70 MB dense blocks, K1 ms per 10 M-query step):
   * t = 9: 155, 10: 141, 11: 112, 12: 107 */
  const uint64_t budget = (ix->d_dense ? 4 : 1) * ix->dev.nblocks * (uint64_t)RS_BLOCK_BYTES;
  uint32_t t = 0;
  while (t < 12 && (16ull << (2 * (t + 1))) <= budget) ++t;
  return t;
}

static int init_streams(rs_index* ix) {
  CU(cudaSetDevice(ix->device));
  CU(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&ix->s_h2d, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&ix->s_d2h, cudaStreamNonBlocking));
  ix->s_k[0] = ix->stream;
  for (int i = 1; i < RS_SEARCH_STREAMS; ++i) CU(cudaStreamCreateWithFlags(&ix->s_k[i], cudaStreamNonBlocking));
  for (int i = 0; i < RS_MAX_CHUNKS; ++i) {
    CU(cudaEventCreateWithFlags(&ix->ev_in[i], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ix->ev_done[i], cudaEventDisableTiming));
  }
  return RS_STATUS_OK;
}

/* host-built arrays -> device (the encoder of rs_build.cpp: kept as the cross-check of rs_encode.cu) */
static int upload(rs_index* ix, const RsHostIndex& h) {
  const int rc = init_streams(ix);
  if (rc != RS_STATUS_OK) return rc;
  const size_t bb = h.blocks.size() * sizeof(rs_block), db = h.dir.size() * sizeof(uint64_t);
  CU(cudaMalloc(&ix->d_blocks, bb));
  CU(cudaMalloc(&ix->d_dir, db));
  CU(cudaMemcpy(ix->d_blocks, h.blocks.data(), bb, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(ix->d_dir, h.dir.data(), db, cudaMemcpyHostToDevice));
  ix->hbm_bytes = bb + db;
  ix->ndir = h.dir.size();
  rs_fill_view(h, &ix->dev);
  for (int c = 0; c < RS_NSYM; ++c) {
    const size_t sb = h.sel[c].size() * sizeof(uint32_t);
    CU(cudaMalloc(&ix->d_sel[c], sb));
    CU(cudaMemcpy(ix->d_sel[c], h.sel[c].data(), sb, cudaMemcpyHostToDevice));
    ix->dev.sel[c] = (const uint32_t*)ix->d_sel[c];
    ix->nsel[c] = h.sel[c].size();
    ix->hbm_bytes += sb;
    ix->init_occ_ref[c] = h.init_occ[c];
  }
  ix->dev.blocks = ix->d_blocks;
  ix->dev.dir = (const uint64_t*)ix->d_dir;
  ix->num_runs = h.num_runs;
  ix->depth = h.depth;
  ix->q2_safe = h.q2_safe;
  if (auto_dense(ix)) {
    const int rc = build_dense(ix);
    if (rc != RS_STATUS_OK) return rc;
  }
  return build_ftab(ix, auto_ftab_k(ix));
}

/* arrays built on the device (rs_encode.cu) become the index's own */
static void adopt(rs_index* ix, rs_encoded& e, uint64_t num_strings, uint64_t num_runs) {
  memset(&ix->dev, 0, sizeof(ix->dev));
  ix->d_blocks = e.d_blocks;
  ix->d_dir = e.d_dir;
  ix->ndir = e.ndir;
  ix->hbm_bytes = e.nblocks * sizeof(rs_block) + e.ndir * 8;
  ix->dev.blocks = e.d_blocks;
  ix->dev.dir = (const uint64_t*)e.d_dir;
  ix->dev.n = e.tail.n;
  ix->dev.nblocks = e.nblocks;
  ix->dev.num_strings = num_strings;
  uint64_t pc = 0;
  for (int c = 0; c < RS_NSYM; ++c) {
    ix->d_sel[c] = e.d_sel[c];
    ix->nsel[c] = e.nsel[c];
    ix->dev.sel[c] = (const uint32_t*)e.d_sel[c];
    ix->dev.sel_shift[c] = e.sel_shift[c];
    ix->hbm_bytes += e.nsel[c] * 4;
    ix->dev.pc[c] = pc;
    pc += e.tail.tot[c];
    ix->dev.tot[c] = e.tail.tot[c];
    ix->init_occ_ref[c] = e.tail.init_occ[c];
  }
  ix->dev.clamp_c = e.tail.clamp_c;
  ix->dev.clamp_char = e.tail.clamp_char;
  ix->dev.vlast = e.tail.vlast;
  ix->num_runs = num_runs;
  ix->depth = e.depth;
  ix->q2_safe = e.tail.q2_safe != 0;
  rs_view_set_tail(ix->init_occ_ref, true, &ix->dev);
  memset(&e, 0, sizeof(e)); /* ownership moved */
}

static int device_ok(int device) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(RS_STATUS_CUDA, "CUDA device %d not available (%d devices): this engine has no CPU path", device, ndev);
  return RS_STATUS_OK;
}

int rs_index_from_runs_dev(const void* d_runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols, int device,
                           rs_index** out) {
  if (!out) return fail(RS_STATUS_ARG, "out is null");
  *out = nullptr;
  int rc = device_ok(device);
  if (rc != RS_STATUS_OK) return rc;
  if (!d_runs || !num_runs) return fail(RS_STATUS_ARG, "no run bytes");
  rs_index* ix = new (std::nothrow) rs_index;
  if (!ix) return fail(RS_STATUS_ARG, "out of host memory");
  ix->device = device;
  rc = init_streams(ix);
  if (rc == RS_STATUS_OK) {
    rs_encoded enc;
    const int erc = rs_encode_device((const uint8_t*)d_runs, num_runs, num_symbols, ix->stream, &enc);
    if (erc == RS_ERR_CUDA) rc = RS_STATUS_CUDA; /* message already recorded */
    else if (erc != RS_OK) rc = fail(erc, "cannot build index from run bytes (status %d)", erc);
    else {
      adopt(ix, enc, num_strings, num_runs);
      rc = auto_dense(ix) ? build_dense(ix) : RS_STATUS_OK;
      if (rc == RS_STATUS_OK) rc = build_ftab(ix, auto_ftab_k(ix));
    }
  }
  if (rc != RS_STATUS_OK) {
    rs_index_close(ix);
    return rc;
  }
  *out = ix;
  return RS_STATUS_OK;
}

int rs_index_from_runs(const uint8_t* runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols, int device,
                       rs_index** out) {
  if (!out) return fail(RS_STATUS_ARG, "out is null");
  *out = nullptr;
  const int rc = device_ok(device);
  if (rc != RS_STATUS_OK) return rc;
  if (!runs || !num_runs) return fail(RS_STATUS_ARG, "no run bytes");
  CU(cudaSetDevice(device));
  uint8_t* d_runs = nullptr;
  CU(cudaMalloc((void**)&d_runs, num_runs));
  cudaError_t e = cudaMemcpy(d_runs, runs, num_runs, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(d_runs);
    return fail(RS_STATUS_CUDA, "copy of the run bytes: %s", cudaGetErrorString(e));
  }
  const int st = rs_index_from_runs_dev(d_runs, num_runs, num_strings, num_symbols, device, out);
  cudaFree(d_runs);
  return st;
}

/* The same index through the HOST encoder (rs_build.cpp, one streaming pass on one thread): the cross-check of
 * the device encoder; tests compare the arrays of the two byte for byte. */
int rs_index_from_runs_host(const uint8_t* runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols, int device,
                            rs_index** out) {
  if (!out) return fail(RS_STATUS_ARG, "out is null");
  *out = nullptr;
  const int drc = device_ok(device);
  if (drc != RS_STATUS_OK) return drc;
  RsHostIndex h;
  int rc;
  try {
    rc = rs_build_host_index(runs, num_runs, num_strings, num_symbols, h);
  } catch (const std::bad_alloc&) {
    return fail(RS_STATUS_ARG, "out of host memory");
  }
  if (rc != RS_OK) return fail(rc, "cannot build index from run bytes (status %d)", rc);
  rs_index* ix = new rs_index;
  ix->device = device;
  const int st = upload(ix, h);
  if (st != RS_STATUS_OK) {
    rs_index_close(ix);
    return st;
  }
  *out = ix;
  return RS_STATUS_OK;
}

int rs_index_open(const char* bwt_path, int device, rs_index** out) {
  if (!bwt_path || !out) return fail(RS_STATUS_ARG, "null argument");
  *out = nullptr;
  std::vector<uint8_t> runs;
  uint64_t ns = 0, nsym = 0;
  int rc;
  try {
    rc = rs_read_bwt_file(bwt_path, runs, ns, nsym);
  } catch (const std::exception&) { /* bad_alloc / length_error on a corrupt header */
    rc = RS_ERR_SHORT;
  }
  if (rc != RS_OK) return fail(rc, "%s: cannot read SGA .bwt (status %d)", bwt_path, rc);
  return rs_index_from_runs(runs.data(), runs.size(), ns, nsym, device, out);
}

/* ---- persisted HBM-layout index (the .bpi2 analogue: rlebwt.cpp:22-31, 150-176; src/util/index_rlebwt.cpp) ---- */
namespace {
struct IndexFileHeader {
  char magic[8]; /* "RSB2IDX1" */
  uint64_t n, num_strings, num_runs, nblocks, ndir;
  uint64_t nsel[RS_NSYM], sel_shift[RS_NSYM], pc[RS_NSYM], tot[RS_NSYM], init_occ_ref[RS_NSYM];
  uint64_t clamp_c, clamp_char, vlast, depth, q2_safe, has_sai;
};
const size_t kIoChunk = 64u << 20;

int write_dev(FILE* f, const void* d, size_t bytes, uint8_t* stage) {
  for (size_t o = 0; o < bytes; o += kIoChunk) {
    const size_t m = bytes - o < kIoChunk ? bytes - o : kIoChunk;
    if (cudaMemcpy(stage, (const uint8_t*)d + o, m, cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
    if (fwrite(stage, 1, m, f) != m) return 2;
  }
  return 0;
}

/* file -> device through two pinned buffers: the read of a chunk overlaps the copy of the previous one */
int read_dev(FILE* f, void* d, size_t bytes, uint8_t* stage[2], cudaEvent_t ev[2], cudaStream_t st, int* turn) {
  for (size_t o = 0; o < bytes; o += kIoChunk) {
    const size_t m = bytes - o < kIoChunk ? bytes - o : kIoChunk;
    const int t = (*turn)++ & 1;
    if (cudaEventSynchronize(ev[t]) != cudaSuccess) return 1;
    if (fread(stage[t], 1, m, f) != m) return 2;
    if (cudaMemcpyAsync((uint8_t*)d + o, stage[t], m, cudaMemcpyHostToDevice, st) != cudaSuccess) return 1;
    if (cudaEventRecord(ev[t], st) != cudaSuccess) return 1;
  }
  return 0;
}
}  // namespace

int rs_index_save(const rs_index* ix, const char* path) {
  if (!ix || !path) return fail(RS_STATUS_ARG, "null argument");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(cudaStreamSynchronize(ix->stream));
  FILE* f = fopen(path, "wb");
  if (!f) return fail(RS_STATUS_OPEN, "cannot create %s", path);
  IndexFileHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "RSB2IDX1", 8);
  h.n = ix->dev.n;
  h.num_strings = ix->dev.num_strings;
  h.num_runs = ix->num_runs;
  h.nblocks = ix->dev.nblocks;
  h.ndir = ix->ndir;
  for (int c = 0; c < RS_NSYM; ++c) {
    h.nsel[c] = ix->nsel[c];
    h.sel_shift[c] = ix->dev.sel_shift[c];
    h.pc[c] = ix->dev.pc[c];
    h.tot[c] = ix->dev.tot[c];
    h.init_occ_ref[c] = ix->init_occ_ref[c];
  }
  h.clamp_c = ix->dev.clamp_c;
  h.clamp_char = ix->dev.clamp_char;
  h.vlast = ix->dev.vlast;
  h.depth = ix->depth;
  h.q2_safe = ix->q2_safe ? 1 : 0;
  h.has_sai = ix->d_sai ? 1 : 0;
  uint8_t* stage = nullptr;
  if (cudaMallocHost((void**)&stage, kIoChunk) != cudaSuccess) {
    fclose(f);
    return fail(RS_STATUS_CUDA, "cannot allocate the staging buffer");
  }
  int bad = fwrite(&h, sizeof(h), 1, f) != 1 ? 2 : 0;
  if (!bad) bad = write_dev(f, ix->d_blocks, h.nblocks * sizeof(rs_block), stage);
  if (!bad) bad = write_dev(f, ix->d_dir, h.ndir * 8, stage);
  for (int c = 0; c < RS_NSYM && !bad; ++c) bad = write_dev(f, ix->d_sel[c], h.nsel[c] * 4, stage);
  if (!bad && ix->d_sai) bad = write_dev(f, ix->d_sai, h.num_strings * 8, stage);
  cudaFreeHost(stage);
  if (fclose(f) != 0 && !bad) bad = 2;
  if (bad == 1) return fail(RS_STATUS_CUDA, "device read failed while saving %s", path);
  if (bad) return fail(RS_STATUS_OPEN, "write to %s failed", path);
  return RS_STATUS_OK;
}

int rs_index_open_encoded(const char* path, int device, rs_index** out) {
  if (!path || !out) return fail(RS_STATUS_ARG, "null argument");
  *out = nullptr;
  int rc = device_ok(device);
  if (rc != RS_STATUS_OK) return rc;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(RS_STATUS_OPEN, "cannot open %s", path);
  IndexFileHeader h;
  if (fread(&h, sizeof(h), 1, f) != 1) {
    fclose(f);
    return fail(RS_STATUS_SHORT, "%s: truncated header", path);
  }
  if (memcmp(h.magic, "RSB2IDX1", 8) != 0) {
    fclose(f);
    return fail(RS_STATUS_MAGIC, "%s is not an encoded index of this library", path);
  }
  /* sizes must agree with the file before anything is allocated from them */
  uint64_t need = sizeof(h) + h.nblocks * sizeof(rs_block) + h.ndir * 8 + (h.has_sai ? h.num_strings * 8 : 0);
  for (int c = 0; c < RS_NSYM; ++c) need += h.nsel[c] * 4;
  fseek(f, 0, SEEK_END);
  const uint64_t have = (uint64_t)ftell(f);
  fseek(f, (long)sizeof(h), SEEK_SET);
  if (have < need || h.nblocks == 0 || h.nblocks > 0xFFFFFFFFull || h.ndir != (h.n >> RS_SLOT_SHIFT) + 2) {
    fclose(f);
    return fail(RS_STATUS_SHORT, "%s: truncated or inconsistent (%llu bytes, header asks for %llu)", path,
                (unsigned long long)have, (unsigned long long)need);
  }
  rs_index* ix = new (std::nothrow) rs_index;
  if (!ix) {
    fclose(f);
    return fail(RS_STATUS_ARG, "out of host memory");
  }
  ix->device = device;
  rc = init_streams(ix);
  uint8_t* stage[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  int turn = 0, bad = 0;
  if (rc == RS_STATUS_OK) {
    cudaError_t e = cudaMallocHost((void**)&stage[0], kIoChunk);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&stage[1], kIoChunk);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc(&ix->d_blocks, h.nblocks * sizeof(rs_block));
    if (e == cudaSuccess) e = cudaMalloc(&ix->d_dir, h.ndir * 8);
    for (int c = 0; c < RS_NSYM && e == cudaSuccess; ++c) e = cudaMalloc(&ix->d_sel[c], (h.nsel[c] ? h.nsel[c] : 1) * 4);
    if (e == cudaSuccess && h.has_sai) e = cudaMalloc((void**)&ix->d_sai, (h.num_strings ? h.num_strings : 1) * 8);
    if (e != cudaSuccess) rc = fail(RS_STATUS_CUDA, "%s: %s", path, cudaGetErrorString(e));
  }
  if (rc == RS_STATUS_OK) {
    cudaStream_t st = ix->s_h2d;
    bad = read_dev(f, ix->d_blocks, h.nblocks * sizeof(rs_block), stage, ev, st, &turn);
    if (!bad) bad = read_dev(f, ix->d_dir, h.ndir * 8, stage, ev, st, &turn);
    for (int c = 0; c < RS_NSYM && !bad; ++c) bad = read_dev(f, ix->d_sel[c], h.nsel[c] * 4, stage, ev, st, &turn);
    if (!bad && h.has_sai) bad = read_dev(f, ix->d_sai, h.num_strings * 8, stage, ev, st, &turn);
    if (cudaStreamSynchronize(st) != cudaSuccess) bad = 1;
    if (bad) rc = fail(bad == 1 ? RS_STATUS_CUDA : RS_STATUS_SHORT, "%s: read failed", path);
  }
  fclose(f);
  for (int t = 0; t < 2; ++t) {
    if (stage[t]) cudaFreeHost(stage[t]);
    if (ev[t]) cudaEventDestroy(ev[t]);
  }
  if (rc == RS_STATUS_OK) {
    memset(&ix->dev, 0, sizeof(ix->dev));
    ix->dev.blocks = ix->d_blocks;
    ix->dev.dir = (const uint64_t*)ix->d_dir;
    ix->dev.n = h.n;
    ix->dev.nblocks = h.nblocks;
    ix->dev.num_strings = h.num_strings;
    ix->ndir = h.ndir;
    ix->hbm_bytes = h.nblocks * sizeof(rs_block) + h.ndir * 8 + (h.has_sai ? h.num_strings * 8 : 0);
    for (int c = 0; c < RS_NSYM; ++c) {
      ix->dev.sel[c] = (const uint32_t*)ix->d_sel[c];
      ix->dev.sel_shift[c] = (uint32_t)h.sel_shift[c];
      ix->nsel[c] = h.nsel[c];
      ix->hbm_bytes += h.nsel[c] * 4;
      ix->dev.pc[c] = h.pc[c];
      ix->dev.tot[c] = h.tot[c];
      ix->init_occ_ref[c] = h.init_occ_ref[c];
    }
    ix->dev.clamp_c = h.clamp_c;
    ix->dev.clamp_char = h.clamp_char;
    ix->dev.vlast = h.vlast;
    ix->num_runs = h.num_runs;
    ix->depth = h.depth;
    ix->q2_safe = h.q2_safe != 0;
    rs_view_set_tail(ix->init_occ_ref, true, &ix->dev);
    rc = auto_dense(ix) ? build_dense(ix) : RS_STATUS_OK;
    if (rc == RS_STATUS_OK) rc = build_ftab(ix, auto_ftab_k(ix));
  }
  if (rc != RS_STATUS_OK) {
    rs_index_close(ix);
    return rc;
  }
  *out = ix;
  return RS_STATUS_OK;
}

/* Copies one of the index's device arrays to the host (tests: device encoder == host encoder; saved == loaded).
 * what: 0 blocks, 1 position directory, 2..6 select directory of $,A,C,G,T, 7 suffix table.  *bytes receives the
 * array's size; the copy is made when buf is non-null and capacity suffices. */
int rs_index_export(const rs_index* ix, int what, void* buf, uint64_t capacity, uint64_t* bytes) {
  if (!ix || !bytes) return fail(RS_STATUS_ARG, "null argument");
  const void* d = nullptr;
  uint64_t n = 0;
  if (what == 0) { d = ix->d_blocks; n = ix->dev.nblocks * sizeof(rs_block); }
  else if (what == 1) { d = ix->d_dir; n = ix->ndir * 8; }
  else if (what >= 2 && what <= 6) { d = ix->d_sel[what - 2]; n = ix->nsel[what - 2] * 4; }
  else if (what == 7) { d = ix->d_ftab; n = ix->d_ftab ? (16ull << (2 * ix->dev.ftab_k)) : 0; }
  else if (what == 8) { d = ix->d_dense; n = ix->dense_bytes; }
  else return fail(RS_STATUS_ARG, "unknown array %d", what);
  *bytes = n;
  if (!buf) return RS_STATUS_OK;
  if (capacity < n) return fail(RS_STATUS_CAPACITY, "array needs %llu bytes", (unsigned long long)n);
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(cudaStreamSynchronize(ix->stream));
  if (n) CU(cudaMemcpy(buf, d, n, cudaMemcpyDeviceToHost));
  return RS_STATUS_OK;
}

void rs_index_close(rs_index* ix) {
  if (!ix) return;
  cudaSetDevice(ix->device);
  if (ix->stream) cudaStreamSynchronize(ix->stream);
  if (ix->d_blocks) cudaFree(ix->d_blocks);
  if (ix->d_dir) cudaFree(ix->d_dir);
  for (int c = 0; c < RS_NSYM; ++c)
    if (ix->d_sel[c]) cudaFree(ix->d_sel[c]);
  if (ix->d_sai) cudaFree(ix->d_sai);
  if (ix->d_samp_off) cudaFree(ix->d_samp_off);
  if (ix->d_samp_codes) cudaFree(ix->d_samp_codes);
  if (ix->d_samp_meta) cudaFree(ix->d_samp_meta);
  if (ix->d_stats) cudaFree(ix->d_stats);
  if (ix->d_ftab) cudaFree(ix->d_ftab);
  if (ix->d_dense) cudaFree(ix->d_dense);
  if (ix->d_loc) cudaFree(ix->d_loc);
  if (ix->d_loc_fin) cudaFree(ix->d_loc_fin);
  if (ix->d_loc_mark) cudaFree(ix->d_loc_mark);
  if (ix->d_loc_rank) cudaFree(ix->d_loc_rank);
  if (ix->d_reads) cudaFree(ix->d_reads);
  if (ix->d_read_off) cudaFree(ix->d_read_off);
  ix->s_in.release();
  ix->s_in2.release();
  ix->s_out.release();
  ix->s_out2.release();
  ix->s_out3.release();
  ix->s_out4.release();
  if (ix->stream) cudaStreamDestroy(ix->stream);
  if (ix->s_h2d) cudaStreamDestroy(ix->s_h2d);
  if (ix->s_d2h) cudaStreamDestroy(ix->s_d2h);
  for (int i = 1; i < RS_SEARCH_STREAMS; ++i)
    if (ix->s_k[i]) cudaStreamDestroy(ix->s_k[i]);
  for (int i = 0; i < RS_MAX_CHUNKS; ++i) {
    if (ix->ev_in[i]) cudaEventDestroy(ix->ev_in[i]);
    if (ix->ev_done[i]) cudaEventDestroy(ix->ev_done[i]);
  }
  delete ix;
}

int rs_index_get_info(const rs_index* ix, rs_index_info* info) {
  if (!ix || !info) return fail(RS_STATUS_ARG, "null argument");
  memset(info, 0, sizeof(*info));
  info->bwlen = ix->dev.n;
  info->num_strings = ix->dev.num_strings;
  info->num_runs = ix->num_runs;
  info->num_blocks = ix->dev.nblocks;
  info->hbm_bytes = ix->hbm_bytes;
  for (int c = 0; c < RS_NSYM; ++c) {
    info->pc[c] = ix->dev.pc[c];
    info->total[c] = ix->dev.tot[c];
  }
  info->ref_depth = ix->depth;
  info->ref_clamp_c = ix->dev.clamp_c;
  info->ref_vlast = ix->dev.vlast;
  info->q2_safe = ix->q2_safe ? 1 : 0;
  info->device = ix->device;
  info->prefix_table_k = (int32_t)ix->dev.ftab_k;
  info->dense_blocks = ix->d_dense ? 1 : 0;
  return RS_STATUS_OK;
}

int rs_index_set_tail_emulation(rs_index* ix, int on) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  std::lock_guard<std::mutex> lk(ix->mu);
  set_tail(ix, on != 0); /* initInterval's constant (query.cpp:18-21) follows the same switch */
  CU(cudaSetDevice(ix->device));
  const int rc = build_ftab(ix, ix->dev.ftab_k); /* the table holds intervals of the semantics it was built with */
  if (rc != RS_STATUS_OK || !ix->d_loc) return rc;
  return build_locate(ix, ix->loc_max_len, ix->loc_every);
}

int rs_index_set_prefix_table(rs_index* ix, int k) {
  if (!ix || k < 0) return fail(RS_STATUS_ARG, "bad argument");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  return build_ftab(ix, (uint32_t)k);
}

int rs_index_enable_stats(rs_index* ix, int on) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  if (on && !ix->d_stats) {
    CU(cudaMalloc((void**)&ix->d_stats, 4 * sizeof(uint64_t)));
    CU(cudaMemset(ix->d_stats, 0, 4 * sizeof(uint64_t)));
  } else if (!on && ix->d_stats) {
    CU(cudaDeviceSynchronize());
    cudaFree(ix->d_stats);
    ix->d_stats = nullptr;
  }
  return RS_STATUS_OK;
}

int rs_index_read_stats(rs_index* ix, uint64_t* interval_steps, uint64_t* walk_steps, int reset) {
  if (!ix) return fail(RS_STATUS_ARG, "null index");
  if (!ix->d_stats) return fail(RS_STATUS_NO_TABLE, "statistics are not enabled");
  CU(cudaSetDevice(ix->device));
  CU(cudaDeviceSynchronize());
  uint64_t v[4] = {0, 0, 0, 0};
  CU(cudaMemcpy(v, ix->d_stats, sizeof(v), cudaMemcpyDeviceToHost));
  if (interval_steps) *interval_steps = v[0];
  if (walk_steps) *walk_steps = v[1];
  if (reset) CU(cudaMemset(ix->d_stats, 0, sizeof(v)));
  return RS_STATUS_OK;
}

int rs_host_alloc(void** p, size_t bytes) {
  if (!p) return fail(RS_STATUS_ARG, "null argument");
  CU(cudaMallocHost(p, bytes ? bytes : 1));
  return RS_STATUS_OK;
}

void rs_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

static inline uint32_t code_of(char b) { return b == 'A' ? 1u : b == 'C' ? 2u : b == 'G' ? 3u : b == 'T' ? 4u : 0u; }

uint64_t rs_get_pc(const rs_index* ix, char b) { return ix ? ix->dev.pc[code_of(b)] : 0; }
uint64_t rs_get_bwlen(const rs_index* ix) { return ix ? ix->dev.n : 0; }

/* ---- host-pointer calls: stage in, launch, stage out on the index's own stream ---- */
int rs_get_occ_batch(const rs_index* ix, const char* b, const uint64_t* index, size_t n, uint64_t* out) {
  if (!ix || (n && (!b || !index || !out))) return fail(RS_STATUS_ARG, "null argument");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(ix->s_in.need(n));
  CU(ix->s_in2.need(n * 8));
  CU(ix->s_out.need(n * 8));
  cudaStream_t st = ix->stream;
  CU(cudaMemcpyAsync(ix->s_in.p, b, n, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ix->s_in2.p, index, n * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_occ(ix->dev, (const uint8_t*)ix->s_in.p, (const uint64_t*)ix->s_in2.p, n, (uint64_t*)ix->s_out.p, st));
  CU(cudaMemcpyAsync(out, ix->s_out.p, n * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

static int char_like(const rs_index* ix, const uint64_t* index, size_t n, char* out, bool f) {
  if (!ix || (n && (!index || !out))) return fail(RS_STATUS_ARG, "null argument");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(ix->s_in.need(n * 8));
  CU(ix->s_out.need(n));
  cudaStream_t st = ix->stream;
  CU(cudaMemcpyAsync(ix->s_in.p, index, n * 8, cudaMemcpyHostToDevice, st));
  if (f) CU(rs_launch_getf(ix->dev, (const uint64_t*)ix->s_in.p, n, (uint8_t*)ix->s_out.p, st));
  else CU(rs_launch_char(ix->dev, (const uint64_t*)ix->s_in.p, n, (uint8_t*)ix->s_out.p, st));
  CU(cudaMemcpyAsync(out, ix->s_out.p, n, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

int rs_get_char_batch(const rs_index* ix, const uint64_t* index, size_t n, char* out) {
  return char_like(ix, index, n, out, false);
}
int rs_get_f_batch(const rs_index* ix, const uint64_t* index, size_t n, char* out) {
  return char_like(ix, index, n, out, true);
}

/* search streams the pipelined calls rotate over (RS_B200_KSTREAMS in the environment overrides) */
static size_t search_streams() {
  static size_t v = 0;
  if (!v) {
    const char* e = getenv("RS_B200_KSTREAMS");
    const long long x = e ? atoll(e) : 0;
    v = (x >= 1 && x <= RS_SEARCH_STREAMS) ? (size_t)x : RS_SEARCH_STREAMS;
  }
  return v;
}

/* queries per pipelined chunk, at least (RS_B200_CHUNK in the environment overrides) */
/* Chunk schedule of the pipelined host-pointer calls.  Every copy costs a few microseconds of its own (1 M
 * 31-mers in and their intervals out, alone on the link: 618 us as one copy each way, 778 us as 15, 875 us as
 * 30), so chunks should be large; but nothing overlaps the first chunk's copy-in nor the last chunk's search and
 * copy-out, so those should be small.  Hence: RS_CHUNK_SMALL, doubling up to the big size, the bulk in equal big
 * chunks, then two small ones.  RS_B200_CHUNK = big size, RS_B200_CHUNK_SMALL = first / last size. */
static size_t env_size(const char* name, size_t dflt) {
  const char* e = getenv(name);
  const long long x = e ? atoll(e) : 0;
  return x >= 1024 ? (size_t)x : dflt;
}
static size_t chunk_big() {
  static size_t v = 0;
  if (!v) v = env_size("RS_B200_CHUNK", 200000);
  return v;
}
static size_t chunk_small() {
  static size_t v = 0;
  if (!v) v = env_size("RS_B200_CHUNK_SMALL", 49152);
  return v;
}
static std::vector<size_t> chunk_schedule(size_t Q) {
  const size_t big = chunk_big(), small = chunk_small() < big ? chunk_small() : big;
  std::vector<size_t> sz;
  const size_t tail = 3 * small; /* 2 * small, small */
  size_t used = 0;
  /* the ramp leaves room for at least one bulk chunk and the two tail chunks among the RS_MAX_CHUNKS events */
  for (size_t up = small; up < big && used + up + tail + big <= Q && sz.size() + 3 < RS_MAX_CHUNKS; up *= 2) {
    sz.push_back(up);
    used += up;
  }
  size_t mid = Q - used - tail;
  size_t nmid = (mid + big - 1) / big;
  if (sz.size() + nmid + 2 > RS_MAX_CHUNKS) nmid = RS_MAX_CHUNKS - 2 - sz.size();
  for (size_t i = 0; i < nmid; ++i) {
    const size_t per = (mid + (nmid - i) - 1) / (nmid - i);
    sz.push_back(per);
    mid -= per;
  }
  sz.push_back(2 * small);
  sz.push_back(small);
  return sz;
}

static int find_host(const rs_index* ix, const char* text, const uint64_t* offsets, int k, size_t Q, uint64_t* lower,
                     uint64_t* upper, int mode, int strand) {
  if (!ix || (Q && (!text || !lower || (mode == 0 && !upper)))) return fail(RS_STATUS_ARG, "null argument");
  if (!offsets && k < 1) return fail(RS_STATUS_ARG, "k must be >= 1 (empty query: SURVEY.md Q3)");
  if (!Q) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  const size_t tbytes = offsets ? (size_t)offsets[Q] : Q * (size_t)k;
  CU(ix->s_in.need(tbytes ? tbytes : 1));
  if (offsets) CU(ix->s_in2.need((Q + 1) * 8));
  CU(ix->s_out.need(Q * 8));
  if (mode == 0) CU(ix->s_out2.need(Q * 8));
  cudaStream_t st = ix->stream;
  uint8_t* d_text = (uint8_t*)ix->s_in.p;
  uint64_t *d_lo = (uint64_t*)ix->s_out.p, *d_up = (uint64_t*)ix->s_out2.p;
  if (!offsets && Q >= chunk_big() + 4 * chunk_small()) {
    /* Large fixed-k batch: split into chunks and run copy-in, search and copy-out of successive chunks
     * on separate streams, so the PCIe transfers hide behind the kernels (and vice versa).  Searches
     * rotate over four streams: the tail of one chunk's kernel (few queries left, SMs draining) overlaps
     * the next ones.  Measured e2e on configs[1]: 1 stream 963, 2: 1050, 4: 1080 M q/s with equal 64k chunks;
     * the copies alone, cut the same way, allow about 1.2-1.3 G q/s (scripts/pcie_probe.py). */
    const std::vector<size_t> sz = chunk_schedule(Q);
    size_t q0 = 0;
    cudaError_t e = cudaSuccess;
    const char* what = "";
#define PIPE(call)                     \
  if (e == cudaSuccess) {              \
    e = (call);                        \
    if (e != cudaSuccess) what = #call; \
  }
    for (size_t c = 0; c < sz.size() && e == cudaSuccess; ++c) {
      const size_t qn = sz[c];
      cudaStream_t sk = ix->s_k[c % search_streams()];
      PIPE(cudaMemcpyAsync(d_text + q0 * k, text + q0 * k, qn * (size_t)k, cudaMemcpyHostToDevice, ix->s_h2d));
      PIPE(cudaEventRecord(ix->ev_in[c], ix->s_h2d));
      PIPE(cudaStreamWaitEvent(sk, ix->ev_in[c], 0));
      PIPE(rs_launch_find_interval(ix->dev, d_text + q0 * k, nullptr, (uint32_t)k, qn, d_lo + q0, d_up + q0, mode, strand,
                                   nullptr, sk));
      PIPE(cudaEventRecord(ix->ev_done[c], sk));
      PIPE(cudaStreamWaitEvent(ix->s_d2h, ix->ev_done[c], 0));
      PIPE(cudaMemcpyAsync(lower + q0, d_lo + q0, qn * 8, cudaMemcpyDeviceToHost, ix->s_d2h));
      if (mode == 0) PIPE(cudaMemcpyAsync(upper + q0, d_up + q0, qn * 8, cudaMemcpyDeviceToHost, ix->s_d2h));
      q0 += qn;
    }
#undef PIPE
    if (e != cudaSuccess) {
      /* copies and kernels already queued still use the scratch buffers and the caller's arrays: drain all three
       * stream groups before handing control back */
      cudaStreamSynchronize(ix->s_h2d);
      for (int i = 0; i < RS_SEARCH_STREAMS; ++i) cudaStreamSynchronize(ix->s_k[i]);
      cudaStreamSynchronize(ix->s_d2h);
      return fail(RS_STATUS_CUDA, "%s: %s", what, cudaGetErrorString(e));
    }
    CU(cudaStreamSynchronize(ix->s_d2h));
    return RS_STATUS_OK;
  }
  CU(cudaMemcpyAsync(d_text, text, tbytes, cudaMemcpyHostToDevice, st));
  if (offsets) CU(cudaMemcpyAsync(ix->s_in2.p, offsets, (Q + 1) * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_find_interval(ix->dev, d_text, offsets ? (const uint64_t*)ix->s_in2.p : nullptr, (uint32_t)k, Q, d_lo,
                             d_up, mode, strand, nullptr, st));
  CU(cudaMemcpyAsync(lower, d_lo, Q * 8, cudaMemcpyDeviceToHost, st));
  if (mode == 0) CU(cudaMemcpyAsync(upper, d_up, Q * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

int rs_find_interval_batch(const rs_index* ix, const char* kmers, int k, size_t Q, uint64_t* lower, uint64_t* upper) {
  return find_host(ix, kmers, nullptr, k, Q, lower, upper, 0, 0);
}

int rs_find_interval_var_batch(const rs_index* ix, const char* text, const uint64_t* offsets, size_t Q,
                               uint64_t* lower, uint64_t* upper) {
  if (Q && !offsets) return fail(RS_STATUS_ARG, "null offsets");
  return find_host(ix, text, offsets, 0, Q, lower, upper, 0, 0);
}

int rs_count_batch(const rs_index* ix, const char* kmers, int k, size_t Q, int strand, uint64_t* counts) {
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 (forward) or 1 (reverse complement)");
  return find_host(ix, kmers, nullptr, k, Q, counts, nullptr, 1, strand);
}

/* ---- device-pointer calls ---- */
int rs_find_interval_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, void* d_lower, void* d_upper,
                               void* stream) {
  if (!ix || (Q && (!d_kmers || !d_lower || !d_upper))) return fail(RS_STATUS_ARG, "null argument");
  if (k < 1) return fail(RS_STATUS_ARG, "k must be >= 1 (empty query: SURVEY.md Q3)");
  CU(rs_launch_find_interval(ix->dev, (const uint8_t*)d_kmers, nullptr, (uint32_t)k, Q, (uint64_t*)d_lower,
                             (uint64_t*)d_upper, 0, 0, (unsigned long long*)ix->d_stats, (cudaStream_t)stream));
  return RS_STATUS_OK;
}

int rs_find_interval_var_batch_dev(const rs_index* ix, const void* d_text, const void* d_offsets, size_t Q,
                                   void* d_lower, void* d_upper, void* stream) {
  if (!ix || (Q && (!d_text || !d_offsets || !d_lower || !d_upper))) return fail(RS_STATUS_ARG, "null argument");
  CU(rs_launch_find_interval(ix->dev, (const uint8_t*)d_text, (const uint64_t*)d_offsets, 0, Q, (uint64_t*)d_lower,
                             (uint64_t*)d_upper, 0, 0, (unsigned long long*)ix->d_stats, (cudaStream_t)stream));
  return RS_STATUS_OK;
}

int rs_count_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, int strand, void* d_counts,
                       void* stream) {
  if (!ix || (Q && (!d_kmers || !d_counts))) return fail(RS_STATUS_ARG, "null argument");
  if (k < 1) return fail(RS_STATUS_ARG, "k must be >= 1");
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 or 1");
  CU(rs_launch_find_interval(ix->dev, (const uint8_t*)d_kmers, nullptr, (uint32_t)k, Q, (uint64_t*)d_counts, nullptr, 1,
                             strand, (unsigned long long*)ix->d_stats, (cudaStream_t)stream));
  return RS_STATUS_OK;
}

/* rs_count_batch_dev without any allocation (no counterpart in the reference; see include/readserver_b200.h) */
int rs_count_batch_noalloc_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, int strand, void* d_counts,
                               void* stream) {
  if (!ix || (Q && (!d_kmers || !d_counts))) return fail(RS_STATUS_ARG, "null argument");
  if (k < 1 || k > 31) return fail(RS_STATUS_ARG, "rs_count_batch_noalloc_dev: k must be in 1..31, got %d", k);
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 or 1");
  if (Q > RS_NOALLOC_MAX_QUERIES)
    return fail(RS_STATUS_ARG, "rs_count_batch_noalloc_dev: %llu queries exceed the static scratch's maximum of %u per call",
                (unsigned long long)Q, (unsigned)RS_NOALLOC_MAX_QUERIES);
  int cur = -1;
  CU(cudaGetDevice(&cur));
  if (cur != ix->device)
    return fail(RS_STATUS_ARG, "rs_count_batch_noalloc_dev: the index lives on device %d, the current device is %d", ix->device, cur);
  CU(rs_launch_count_noalloc(ix->dev, (const uint8_t*)d_kmers, (uint32_t)k, Q, strand, (uint64_t*)d_counts,
                             (unsigned long long*)ix->d_stats, (cudaStream_t)stream));
  return RS_STATUS_OK;
}

/* ---- locate ---- */
int rs_index_set_sai(rs_index* ix, const uint64_t* sai, uint64_t num_strings) {
  if (!ix || !sai) return fail(RS_STATUS_ARG, "null argument");
  if (num_strings != ix->dev.num_strings)
    return fail(RS_STATUS_ARG, "sai has %llu entries, the index holds %llu strings", (unsigned long long)num_strings,
                (unsigned long long)ix->dev.num_strings);
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  if (ix->d_sai) {
    cudaFree(ix->d_sai);
    ix->hbm_bytes -= num_strings * 8;
    ix->d_sai = nullptr;
  }
  CU(cudaMalloc((void**)&ix->d_sai, num_strings ? num_strings * 8 : 8));
  CU(cudaMemcpy(ix->d_sai, sai, num_strings * 8, cudaMemcpyHostToDevice));
  ix->hbm_bytes += num_strings * 8;
  return RS_STATUS_OK;
}

int rs_index_set_sai_dev(rs_index* ix, const void* d_sai, uint64_t num_strings) {
  if (!ix || !d_sai) return fail(RS_STATUS_ARG, "null argument");
  if (num_strings != ix->dev.num_strings)
    return fail(RS_STATUS_ARG, "sai has %llu entries, the index holds %llu strings", (unsigned long long)num_strings,
                (unsigned long long)ix->dev.num_strings);
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  if (ix->d_sai) {
    cudaFree(ix->d_sai);
    ix->hbm_bytes -= num_strings * 8;
    ix->d_sai = nullptr;
  }
  CU(cudaMalloc((void**)&ix->d_sai, num_strings ? num_strings * 8 : 8));
  CU(cudaMemcpy(ix->d_sai, d_sai, num_strings * 8, cudaMemcpyDeviceToDevice));
  ix->hbm_bytes += num_strings * 8;
  return RS_STATUS_OK;
}

int rs_extract_prefix_batch_dev(const rs_index* ix, const void* d_rows, size_t n, uint32_t max_steps, void* d_lexrank,
                                void* d_steps, void* d_read_index, void* d_bases, void* stream) {
  if (!ix || (n && !d_rows)) return fail(RS_STATUS_ARG, "null argument");
  if (d_read_index && !ix->d_sai) return fail(RS_STATUS_NO_TABLE, "read_index requested but no .sai table attached");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  CU(rs_launch_extract_prefix(ix->dev, (const uint64_t*)d_rows, n, max_steps, (uint64_t*)d_lexrank, (uint32_t*)d_steps,
                              ix->d_sai, (uint64_t*)d_read_index, (uint8_t*)d_bases, (unsigned long long*)ix->d_stats,
                              (cudaStream_t)stream));
  return RS_STATUS_OK;
}

int rs_extract_prefix_batch(const rs_index* ix, const uint64_t* rows, size_t n, uint32_t max_steps, uint64_t* lexrank,
                            uint32_t* steps, uint64_t* read_index, char* bases) {
  if (!ix || (n && !rows)) return fail(RS_STATUS_ARG, "null argument");
  if (read_index && !ix->d_sai) return fail(RS_STATUS_NO_TABLE, "read_index requested but no .sai table attached");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(ix->s_in.need(n * 8));
  CU(ix->s_out.need(n * 8));
  CU(ix->s_out2.need(n * 4));
  if (read_index) CU(ix->s_out3.need(n * 8));
  if (bases) CU(ix->s_out4.need(n * (size_t)max_steps));
  cudaStream_t st = ix->stream;
  CU(cudaMemcpyAsync(ix->s_in.p, rows, n * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_extract_prefix(ix->dev, (const uint64_t*)ix->s_in.p, n, max_steps, (uint64_t*)ix->s_out.p,
                              (uint32_t*)ix->s_out2.p, ix->d_sai, read_index ? (uint64_t*)ix->s_out3.p : nullptr,
                              bases ? (uint8_t*)ix->s_out4.p : nullptr, nullptr, st));
  if (lexrank) CU(cudaMemcpyAsync(lexrank, ix->s_out.p, n * 8, cudaMemcpyDeviceToHost, st));
  if (steps) CU(cudaMemcpyAsync(steps, ix->s_out2.p, n * 4, cudaMemcpyDeviceToHost, st));
  if (read_index) CU(cudaMemcpyAsync(read_index, ix->s_out3.p, n * 8, cudaMemcpyDeviceToHost, st));
  if (bases) CU(cudaMemcpyAsync(bases, ix->s_out4.p, n * (size_t)max_steps, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

/* ---- select / extractPostfix ---- */
int rs_get_occ_at_batch(const rs_index* ix, const char* b, const uint64_t* count, size_t n, uint64_t* out) {
  if (!ix || (n && (!b || !count || !out))) return fail(RS_STATUS_ARG, "null argument");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(ix->s_in.need(n));
  CU(ix->s_in2.need(n * 8));
  CU(ix->s_out.need(n * 8));
  cudaStream_t st = ix->stream;
  CU(cudaMemcpyAsync(ix->s_in.p, b, n, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ix->s_in2.p, count, n * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_occ_at(ix->dev, (const uint8_t*)ix->s_in.p, (const uint64_t*)ix->s_in2.p, n, (uint64_t*)ix->s_out.p, st));
  CU(cudaMemcpyAsync(out, ix->s_out.p, n * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}

int rs_extract_postfix_batch_dev(const rs_index* ix, const void* d_rows, size_t n, uint32_t max_steps,
                                 void* d_read_index, void* d_steps, void* d_bases, void* stream) {
  if (!ix || (n && !d_rows)) return fail(RS_STATUS_ARG, "null argument");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  CU(rs_launch_extract_postfix(ix->dev, (const uint64_t*)d_rows, n, max_steps, (uint64_t*)d_read_index,
                               (uint32_t*)d_steps, (uint8_t*)d_bases, (cudaStream_t)stream));
  return RS_STATUS_OK;
}

int rs_extract_postfix_batch(const rs_index* ix, const uint64_t* rows, size_t n, uint32_t max_steps,
                             uint64_t* read_index, uint32_t* steps, char* bases) {
  if (!ix || (n && !rows)) return fail(RS_STATUS_ARG, "null argument");
  if (max_steps < 1) return fail(RS_STATUS_ARG, "max_steps must be >= 1");
  if (!n) return RS_STATUS_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  CU(cudaSetDevice(ix->device));
  CU(ix->s_in.need(n * 8));
  CU(ix->s_out.need(n * 8));
  CU(ix->s_out2.need(n * 4));
  if (bases) CU(ix->s_out4.need(n * (size_t)max_steps));
  cudaStream_t st = ix->stream;
  CU(cudaMemcpyAsync(ix->s_in.p, rows, n * 8, cudaMemcpyHostToDevice, st));
  CU(rs_launch_extract_postfix(ix->dev, (const uint64_t*)ix->s_in.p, n, max_steps, (uint64_t*)ix->s_out.p,
                               (uint32_t*)ix->s_out2.p, bases ? (uint8_t*)ix->s_out4.p : nullptr, st));
  if (read_index) CU(cudaMemcpyAsync(read_index, ix->s_out.p, n * 8, cudaMemcpyDeviceToHost, st));
  if (steps) CU(cudaMemcpyAsync(steps, ix->s_out2.p, n * 4, cudaMemcpyDeviceToHost, st));
  if (bases) CU(cudaMemcpyAsync(bases, ix->s_out4.p, n * (size_t)max_steps, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RS_STATUS_OK;
}
