/* Launchers of the sm_100a kernels (rs_kernels.cu); device pointers, stream-ordered. */
#ifndef RS_KERNELS_H
#define RS_KERNELS_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rs_layout.h"

#define RS_THREADS 256

cudaError_t rs_launch_occ(const rs_index_dev& ix, const uint8_t* syms, const uint64_t* idx, uint64_t n, uint64_t* out,
                          cudaStream_t st);
cudaError_t rs_launch_char(const rs_index_dev& ix, const uint64_t* idx, uint64_t n, uint8_t* out, cudaStream_t st);
cudaError_t rs_launch_getf(const rs_index_dev& ix, const uint64_t* idx, uint64_t n, uint8_t* out, cudaStream_t st);
#define RS_MAX_RANKS 16
/* accumulation buffers of the ranks of one box, mapped into this process (rs_peer.cu) */
typedef struct rs_peers {
  uint64_t* buf[RS_MAX_RANKS];
  uint32_t world, rank;
} rs_peers;

/* mode 0: lower/upper; mode 1: counts into `lower` (strand 0 forward, 1 reverse complement);
 * mode 2: counts added into the owning rank's buffer through peer memory (`lower` = Q words of scratch,
 * fixed-length queries of k <= 31 only); mode 3: mode 2 reusing the packed queries an earlier mode-2 call on the
 * same batch left in `lower` (several shards searched for one batch: the pre-pass runs once);
 * stats (nullable): += number of updateInterval steps executed (2 ranks each) */
cudaError_t rs_launch_find_interval(const rs_index_dev& ix, const uint8_t* text, const uint64_t* offs, uint32_t k,
                                    uint64_t Q, uint64_t* lower, uint64_t* upper, int mode, int strand,
                                    unsigned long long* stats, cudaStream_t st, const rs_peers* peers = nullptr);
cudaError_t rs_launch_pack_kmers(const uint8_t* text, uint32_t k, uint64_t Q, int rc, uint64_t* packed, cudaStream_t st);
/* K1 over packed keys kept apart from its outputs (mode 4: first row + row count per key; mode 2 as above) */
cudaError_t rs_launch_find_interval_packed(const rs_index_dev& ix, const uint64_t* pkeys, uint32_t k, uint64_t Q,
                                           uint64_t* first, uint32_t* rows, int mode, unsigned long long* stats,
                                           cudaStream_t st, const rs_peers* peers = nullptr,
                                           unsigned long long* ticket = nullptr);
/* Count (mode 1) with NO allocation: the packed keys and the dense K1's work counter live in a file-scope device array
 * of rs_kernels.cu (RS_NOALLOC_MAX_QUERIES * 8 + 256 bytes per GPU), shared by all calls on the device. */
#ifndef RS_NOALLOC_MAX_QUERIES
#define RS_NOALLOC_MAX_QUERIES (1u << 20) /* = include/readserver_b200.h */
#endif
cudaError_t rs_launch_count_noalloc(const rs_index_dev& ix, const uint8_t* text, uint32_t k, uint64_t Q, int strand,
                                    uint64_t* counts, unsigned long long* stats, cudaStream_t st);
/* work counters of the kernels that hand their items out dynamically: `n` consecutive cleared slots, one per launch */
cudaError_t rs_ticket_take(uint32_t n, unsigned long long** first, cudaStream_t st);
/* cross-GPU barrier over peer-mapped flag words: every rank writes `epoch` into its slot of every peer's
 * flag array, then waits until all slots of its own array reached it (or ~timeout_ms passed: *err = 1) */
cudaError_t rs_launch_peer_barrier(uint64_t* const* peer_flags, uint32_t world, uint32_t rank, uint64_t epoch,
                                   uint32_t timeout_ms, uint32_t* err, cudaStream_t st);
cudaError_t rs_build_ftab(const rs_index_dev& ix, uint32_t t, uint64_t* table, cudaStream_t st);
cudaError_t rs_launch_extract_prefix(const rs_index_dev& ix, const uint64_t* rows, uint64_t n, uint32_t max_steps,
                                     uint64_t* lexrank, uint32_t* steps, const uint64_t* sai, uint64_t* read_index,
                                     uint8_t* bases, unsigned long long* stats, cudaStream_t st);
cudaError_t rs_launch_extract_prefix_n(const rs_index_dev& ix, const uint64_t* rows, const uint64_t* n_dev, uint64_t n,
                                       uint32_t max_steps, const uint64_t* sai, uint64_t* read_index,
                                       unsigned long long* stats, cudaStream_t st);

/* fills loc[n] and fin[num_strings] (rs_layout.h: locate table); walks are cut at max_len symbols */
cudaError_t rs_build_locate_table(const rs_index_dev& ix, uint32_t max_len, uint64_t* loc, uint64_t* fin, cudaStream_t st);

/* sampled locate table: rec = one 16-byte record per 64 rows {marker bits, sampled rows before the word} */
cudaError_t rs_locate_mark(const rs_index_dev& ix, uint32_t max_len, uint32_t every, uint64_t* rec, uint64_t* fin, cudaStream_t st);
cudaError_t rs_locate_mark_counts(const uint64_t* rec, uint64_t n, uint64_t* counts, cudaStream_t st);
cudaError_t rs_locate_mark_ranks(uint64_t* rec, uint64_t n, const uint64_t* before, cudaStream_t st);
cudaError_t rs_locate_fill(const rs_index_dev& ix, uint32_t max_len, uint32_t every, const uint64_t* rec, uint64_t* loc,
                           uint64_t nsamples, cudaStream_t st);

cudaError_t rs_launch_occ_at(const rs_index_dev& ix, const uint8_t* syms, const uint64_t* cnt, uint64_t n, uint64_t* out,
                             cudaStream_t st);
cudaError_t rs_launch_extract_postfix(const rs_index_dev& ix, const uint64_t* rows, uint64_t n, uint32_t max_steps,
                                      uint64_t* read_index, uint32_t* steps, uint8_t* bases, cudaStream_t st);

/* dense rank blocks (rs_layout.h) from the run blocks; *bad |= 1 when a count does not fit 32 bits */
cudaError_t rs_build_dense(const rs_index_dev& ix, uint64_t ndense, void* out, unsigned int* bad, cudaStream_t st);

/* rs_pipeline.cu */
cudaError_t rs_sort_pairs_bytes(uint64_t n, size_t* bytes);
cudaError_t rs_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                          uint64_t n, int end_bit, void* tmp, size_t tmp_bytes, cudaStream_t st);
cudaError_t rs_scan_bytes(uint64_t n, size_t* bytes);
cudaError_t rs_exclusive_scan(const uint64_t* counts, uint64_t* offsets, uint64_t n, void* tmp, size_t tmp_bytes,
                              cudaStream_t st);
cudaError_t rs_launch_hit_counts(const uint64_t* lower, const uint64_t* upper, uint64_t Q, uint64_t max_hits,
                                 uint64_t* counts, cudaStream_t st);
cudaError_t rs_launch_expand_rows(const uint64_t* lower, const uint64_t* offsets, uint64_t Q, uint64_t capacity,
                                  uint64_t* rows, uint32_t* owner, cudaStream_t st);
cudaError_t rs_launch_sample_counts(const uint64_t* samp_off, uint64_t num_strings, const uint64_t* read_index,
                                    uint64_t n, uint64_t* counts, cudaStream_t st);
cudaError_t rs_launch_sample_gather(const uint64_t* samp_off, const uint8_t* samp_codes, uint32_t sos,
                                    uint64_t num_strings, const uint64_t* read_index, uint64_t n,
                                    const uint64_t* out_off, uint8_t* out_codes, cudaStream_t st);
cudaError_t rs_launch_carriers(const uint64_t* samp_off, const uint8_t* samp_codes, uint32_t sos, uint64_t num_strings,
                               uint32_t num_samples, const uint64_t* hit_off, const uint64_t* read_index, uint64_t Q,
                               int emit, uint64_t* counts, const uint64_t* out_off, uint8_t* out_codes,
                               cudaStream_t st);

#endif
