/* Device-side re-encoder (rs_encode.cu): run bytes in HBM -> blocks and directories in HBM. */
#ifndef RS_ENCODE_H
#define RS_ENCODE_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"

typedef struct rs_enc_tail {
  uint64_t n;
  uint64_t tot[RS_NSYM];
  uint64_t init_occ[RS_NSYM]; /* the reference's getOcc(c, n - 1) (initInterval, query.cpp:18-21) */
  uint64_t clamp_c, clamp_char, vlast;
  uint32_t q2_safe, pad_;
} rs_enc_tail;

typedef struct rs_encoded {
  void* d_blocks;
  void* d_dir;
  void* d_sel[RS_NSYM];
  uint64_t nblocks, ndir, nsel[RS_NSYM];
  uint32_t sel_shift[RS_NSYM];
  uint64_t depth;
  rs_enc_tail tail;
} rs_encoded;

/* returns RS_OK or an RS_ERR_* of rs_build.h; on success the caller owns the device arrays */
int rs_encode_device(const uint8_t* d_runs, uint64_t R, uint64_t num_symbols, cudaStream_t st, rs_encoded* out);
void rs_encoded_free(rs_encoded* e);

#endif
