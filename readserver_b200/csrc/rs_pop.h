/* Device-side tables of a shard set (rs_pop.cu), passed to its kernels by value. */
#ifndef RS_POP_H
#define RS_POP_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"

#define RS_POP_MAX_SHARDS 1024
#define RS_POP_MAX_SHARD_ID 4096u
#define RS_POP_SCRATCH_BYTES (12ull << 30) /* per-shard interval scratch of one chunk, at most */
#define RS_POP_MAX_SUB 16 /* sub-batches of a host-pointer call (copy-out overlapped with the next one's kernels) */
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


/* Where the results of a step go.  One GPU: this process's own buffers.  Several GPUs (one process each, rs_popcomm):
 * every rank's buffers as mapped into this process through CUDA IPC (NVLink peer memory), indexed by rank. */
#ifndef RS_MAX_RANKS
#define RS_MAX_RANKS 16
#endif
typedef struct rs_pop_peers {
  uint32_t world, rank;
  uint64_t maxq;                    /* row stride of rankh                                                     */
  uint64_t* counts[RS_MAX_RANKS];   /* [maxq] population counts: the owner's entries are the sums               */
  uint32_t* rankh[RS_MAX_RANKS];    /* [world][maxq] hits per query and source rank, written by the sources     */
  uint64_t* slot[RS_MAX_RANKS];     /* [maxq] where THAT rank's hits of a query start in the owner's hit list   */
  uint64_t* hits[RS_MAX_RANKS];     /* hit lists                                                                */
  uint32_t* carr[RS_MAX_RANKS];     /* [maxq x 4] carrier bitmaps                                               */
} rs_pop_peers;

#endif
