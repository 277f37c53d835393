/* Device-side tables of a shard set (rs_pop.cu), passed to its kernels by value. */
#ifndef RS_POP_H
#define RS_POP_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"

#define RS_POP_MAX_SHARDS 1024
#define RS_POP_MAX_SHARD_ID 4096u
#define RS_POP_STAGES 4
#define RS_POP_STAGE_PACK 0    /* k_pack_kmers                                  */
#define RS_POP_STAGE_SEARCH 1  /* K1, one launch per shard and chunk            */
#define RS_POP_STAGE_OFFSETS 2 /* totals, scan, offsets                         */
#define RS_POP_STAGE_LOCATE 3  /* gather from the tables / expand + walk + scatter, carrier sets */

typedef struct rs_pop_tables {
  uint32_t nshards;
  uint32_t max_steps;                  /* bound of the fallback walks = max_read_len of the seal */
  uint32_t sos;                        /* size_of_sample */
  const uint32_t* const* rid;          /* [nshards] read-index tables (null entries: not built)          */
  const uint4* const* sets;            /* [nshards] per-read sample-set bitmaps, 128 samples (or null)    */
  const uint64_t* const* sai;          /* [nshards] lexicographic rank -> read index                       */
  const uint64_t* const* samp_off;     /* [nshards] sample tables (CSR)                                   */
  const uint8_t* const* samp_codes;
  const uint64_t* tag;                 /* [nshards] shard_id << 40                                         */
  const uint32_t* slot_of_shard;       /* [RS_POP_MAX_SHARD_ID] shard id -> position in the set            */
  const rs_index_dev* views;           /* [nshards] kernel views (for the rare rows the tables do not cover) */
} rs_pop_tables;

#endif
