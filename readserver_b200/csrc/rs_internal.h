/* Internals shared by the C-ABI translation units (not part of the public interface). */
#ifndef RS_INTERNAL_H
#define RS_INTERNAL_H

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/readserver_b200.h"
#include "rs_build.h"
#include "rs_kernels.h"

#define RS_MAX_CHUNKS 32
#define RS_SEARCH_STREAMS 4

int rs_fail(int code, const char* fmt, ...); /* records the thread's last error message */
#define fail rs_fail

#define CU(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) return fail(RS_STATUS_CUDA, "%s: %s", #call, cudaGetErrorString(e_));  \
  } while (0)

/* grow-on-demand device scratch buffer */
struct Scratch {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t need(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

/* stream-ordered temporaries of a multi-kernel call (freed in stream order on scope exit) */
struct Tmp {
  cudaStream_t st;
  std::vector<void*> ptrs;
  explicit Tmp(cudaStream_t s) : st(s) {}
  cudaError_t get(void** p, size_t bytes) {
    cudaError_t e = cudaMallocAsync(p, bytes ? bytes : 8, st);
    if (e == cudaSuccess) ptrs.push_back(*p);
    return e;
  }
  ~Tmp() {
    for (void* p : ptrs) cudaFreeAsync(p, st);
  }
};

struct rs_index {
  int device = 0;
  rs_index_dev dev;
  uint64_t num_runs = 0, depth = 0;
  bool q2_safe = true;
  uint64_t hbm_bytes = 0;
  uint64_t init_occ_ref[RS_NSYM] = {0, 0, 0, 0, 0}; /* the reference's getOcc(c, n-1) */
  void* d_blocks = nullptr;
  void* d_dir = nullptr;
  uint64_t ndir = 0, nsel[RS_NSYM] = {0, 0, 0, 0, 0}; /* entries of the position / select directories */
  void* d_sel[RS_NSYM] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  uint64_t* d_sai = nullptr;
  uint64_t* d_samp_off = nullptr;
  uint8_t* d_samp_codes = nullptr;
  uint8_t* d_samp_meta = nullptr; /* 2 bytes per code (count + 33, lane + 33) when the store has_other_meta_data */
  uint64_t samp_bytes = 0;
  int size_of_sample = 0;
  uint32_t num_samples = 0; /* 1 + largest sample number that occurs in the table */
  uint32_t* d_reads = nullptr;     /* packed read table (rs_reads.cu): 16 bases per word */
  uint64_t* d_read_off = nullptr;  /* first base of each read, num_strings + 1 entries */
  uint64_t reads_bytes = 0;
  uint64_t* d_loc = nullptr;      /* locate table (rs_layout.h), n entries */
  uint64_t* d_loc_fin = nullptr;  /* num_strings entries */
  uint64_t* d_loc_mark = nullptr; /* sampled variant: one bit per row, and its rank directory */
  uint64_t* d_loc_rank = nullptr;
  uint32_t loc_max_len = 0, loc_every = 0;
  uint64_t loc_bytes = 0;
  void* d_dense = nullptr;    /* dense rank blocks (rs_layout.h), optional */
  uint64_t dense_bytes = 0;
  uint64_t* d_ftab = nullptr; /* k-mer suffix table (rs_layout.h), 4^ftab_k pairs */
  uint64_t* d_stats = nullptr; /* [0] updateInterval steps, [1] LF-walk steps; null unless profiling is on */
  cudaStream_t stream = nullptr;
  /* copy-in / copy-out streams and per-chunk events of the pipelined host-pointer calls */
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaStream_t s_k[RS_SEARCH_STREAMS] = {}; /* search streams of the pipelined calls (s_k[0] = stream) */
  cudaEvent_t ev_in[RS_MAX_CHUNKS] = {}, ev_done[RS_MAX_CHUNKS] = {};
  mutable std::mutex mu; /* serialises the host-pointer calls (they share the scratch buffers) */
  mutable Scratch s_in, s_in2, s_out, s_out2, s_out3, s_out4;
};

#endif
