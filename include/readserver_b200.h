/* readserver_b200 -- C ABI of the B200-native FM-index query engine.
 *
 * Drop-in boundary for ReadServer's BWT hot path.  The reference has no FFI: its seam is the C++
 * static library libbwt.a (/root/reference/src/bwt/Makefile:11) with the abstract `class BWT`
 * (include/bwt/bwt.h:6-15) and the free functions of include/bwt/query.h:18-32, consumed by the
 * query worker (src/service/service.cpp).  Each entry point below names the reference interface it
 * replaces.  All calls are batched; plain pointers and sizes only.
 *
 * Conventions
 *  - every function returns an rs_status (0 = OK); rs_last_error() gives the thread's last message;
 *  - symbols are the reference's characters '$','A','C','G','T'; any other byte ranks as '$'
 *    (include/bwt/alphabet.h:11-32);
 *  - results are bit-exact with the reference's CPU path INCLUDING its tail behaviour in
 *    BPTree::rank/access (include/bwt/BPTree.h:74-94,135-157) unless rs_index_set_tail_emulation(ix, 0);
 *  - "_dev" variants take device pointers and a cudaStream_t (as void*), are asynchronous and
 *    stream-ordered; the others take host pointers and return when the results are in them;
 *  - an rs_index is immutable after its tables are attached and may be shared by any number of
 *    threads (the reference shares one `const BWT*` between 8 + 64 threads, service.cpp:88-89,1505-1513);
 *    host-pointer calls on one index are serialised internally, "_dev" calls are not;
 *  - there is NO CPU fallback: without a CUDA device every compute call fails with RS_ERR_CUDA.
 */
#ifndef READSERVER_B200_H
#define READSERVER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rs_index rs_index;

typedef enum rs_status {
  RS_STATUS_OK = 0,
  RS_STATUS_OPEN = 1,      /* cannot open file (the reference silently yields an empty index)          */
  RS_STATUS_MAGIC = 2,     /* bad magic: the reference prints and exit(EXIT_FAILURE)s, rlebwt_reader.cpp:31-34 */
  RS_STATUS_SHORT = 3,     /* truncated file                                                            */
  RS_STATUS_RUNBYTE = 4,   /* run byte with symbol > 4 or length 0                                      */
  RS_STATUS_ARG = 5,       /* null / empty / inconsistent argument (e.g. empty query)     */
  RS_STATUS_CUDA = 6,      /* CUDA runtime error                                                        */
  RS_STATUS_TOO_LARGE = 7, /* more than 2^32 blocks in one shard                                        */
  RS_STATUS_NO_TABLE = 8,  /* locate / sample call before the table it needs was attached              */
  RS_STATUS_CAPACITY = 9   /* an output buffer is too small; the sizes needed have been written          */
} rs_status;

typedef struct rs_index_info {
  uint64_t bwlen;        /* getBWLen(), rlebwt.cpp:303-305                                   */
  uint64_t num_strings;  /* reads in the shard                                                */
  uint64_t num_runs;     /* run bytes of the .bwt                                             */
  uint64_t num_blocks;   /* 64-byte blocks in HBM                                             */
  uint64_t hbm_bytes;    /* blocks + directories + attached tables                            */
  uint64_t pc[5];        /* getPC for $,A,C,G,T                                               */
  uint64_t total[5];     /* occurrences of each symbol                                        */
  uint64_t ref_depth;    /* depth of the BP-tree the reference would build (rlebwt.cpp:51-72) */
  uint64_t ref_clamp_c;  /* first position of the reference's tail region      */
  uint64_t ref_vlast;    /* index of the last m_vSum hint                                     */
  int32_t q2_safe;       /* 0: the reference would read m_vSum out of bounds   */
  int32_t device;        /* CUDA device ordinal                                               */
  int32_t prefix_table_k; /* length of the k-mer suffix table (0 = none)                      */
  int32_t dense_blocks;   /* 1: dense rank blocks are built (rs_index_set_dense)              */
} rs_index_info;

/* ---- lifecycle -------------------------------------------------------------------------------- */
/* RLEBWT::RLEBWT(filename) (src/bwt/rlebwt.cpp:11-32) + BWTReaderRLE (src/bwt/rlebwt_reader.cpp:21-48):
 * read an SGA binary RLE .bwt, re-encode it into 64-byte blocks and upload it to `device`. */
int rs_index_open(const char* bwt_path, int device, rs_index** out);
/* Same from run bytes already in memory (sym<<5 | len, include/bwt/rlunit.h:8-16). */
int rs_index_from_runs(const uint8_t* runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols,
                       int device, rs_index** out);
/* Same from run bytes already in device memory on `device` (no host staging at all).  The re-encoding itself --
 * RLEBWT::initialiseFMIndex, rlebwt.cpp:34-148, one pass on one host thread in the reference -- runs on the GPU
 * for all three (rs_encode.cu). */
int rs_index_from_runs_dev(const void* d_runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols,
                           int device, rs_index** out);
/* The same index built by the host-side encoder (one streaming pass, rs_build.cpp): kept as the cross-check of
 * the device encoder -- tests compare the arrays of the two byte for byte (rs_index_export). */
int rs_index_from_runs_host(const uint8_t* runs, uint64_t num_runs, uint64_t num_strings, uint64_t num_symbols,
                            int device, rs_index** out);
/* Persisted HBM-layout index: what the reference's .bpi2 file is to its BP-tree (RLEBWT::serialiseFMIndex /
 * deserialiseFMIndex, rlebwt.cpp:22-31, 150-200; written by src/util/index_rlebwt.cpp).  rs_index_save writes the
 * blocks, both directories, the constants and -- when attached -- the .sai table; rs_index_open_encoded loads such
 * a file without touching the .bwt (file -> pinned staging -> HBM, read and copy overlapped) and rebuilds only
 * the k-mer suffix table. */
int rs_index_save(const rs_index* ix, const char* path);
int rs_index_open_encoded(const char* path, int device, rs_index** out);
/* Copy of one device array of the index (0 blocks, 1 position directory, 2..6 select directory of $,A,C,G,T,
 * 7 k-mer suffix table, 8 dense rank blocks): *bytes = its size; copied when buf != NULL and capacity suffices. */
int rs_index_export(const rs_index* ix, int what, void* buf, uint64_t capacity, uint64_t* bytes);
void rs_index_close(rs_index* ix);
int rs_index_get_info(const rs_index* ix, rs_index_info* info);
/* 1 (default): bit-exact with the reference; 0: exact rank/access everywhere. */
int rs_index_set_tail_emulation(rs_index* ix, int on);
/* k-mer suffix table, no reference counterpart (results do not depend on it): the intervals of all
 * 4^k strings of length k, so that findInterval starts k-1 steps in.  Built automatically at open
 * (k <= 10, at most a quarter of the blocks' size; RS_B200_FTAB=k overrides); k = 0 drops it. */
int rs_index_set_prefix_table(rs_index* ix, int k);
/* Dense rank blocks, no reference counterpart (results do not depend on them): the BWT once more as 128 symbols per
 * 64-byte block in three bit planes with the four base counts in front -- addressed by position alone, so a rank
 * (getOcc) or access (getChar) is ONE memory access and a few logic + popcount instructions instead of a directory
 * word, a run block and a scan over 36 run bytes.  0.5 bytes per BWT symbol beside the run blocks, which keep
 * serving select (getOccAt) and the reference's rule for bytes outside "$ACGT".  Built automatically at open when
 * they cost at most twice the run blocks -- runs of ~7 symbols or shorter, as in the shards of a read-partitioned
 * population -- and every base occurs fewer than 2^32 times (RS_B200_DENSE = 0 / 1 in the environment: never /
 * whenever possible); on = 0 drops them. */
int rs_index_set_dense(rs_index* ix, int on);
/* Locate table, no reference counterpart (results do not depend on it): for every BWT row, where
 * extractPrefix's walk from it ends and after how many bases, found once by walking every read from its
 * terminator (n LF steps in one kernel).  With it the locate stage (rs_extract_prefix_batch without bases,
 * rs_locate_batch, rs_gt_reads_batch) reads 8 bytes per hit row instead of walking ~read-length/2 blocks.
 * Costs 8 bytes per BWT symbol of HBM (8.7 GB for a 1.08 G-symbol shard): meant for shards that leave the room.
 * max_read_len bounds the walks of the build (rows on longer walks -- the reference's tail region can cycle --
 * stay uncovered and are walked at query time as before).  Rebuilt by rs_index_set_tail_emulation. */
int rs_index_build_locate_table(rs_index* ix, uint32_t max_read_len);
/* The same, sampled: entries only for the rows reached after a multiple of `every` bases of a read's walk (a
 * 16-byte record per 64 rows holds their marker bits and the number of entries before them), so a hit row is walked
 * fewer than `every` steps instead of to the '$'.  n (1/4 + 8/every) bytes: 0.75 bytes per symbol at every = 16.
 * every = 1 is the full table. */
int rs_index_build_locate_samples(rs_index* ix, uint32_t max_read_len, uint32_t every);
int rs_index_drop_locate_table(rs_index* ix);
/* Work counters for the *_dev calls (kept on the device, off by default): executed updateInterval
 * steps of findInterval (2 getOcc each) and executed LF steps of extractPrefix.  Used by bench.py
 * to turn a kernel time into achieved bytes/s; no reference counterpart. */
int rs_index_enable_stats(rs_index* ix, int on);
int rs_index_read_stats(rs_index* ix, uint64_t* interval_steps, uint64_t* walk_steps, int reset);
/* Measurement aid, no reference counterpart: achieved GB/s of random, 128-byte-aligned, 128-byte
 * loads over a buffer_bytes buffer on `device` -- the random-access roofline of this path. */
int rs_measure_random_access(int device, uint64_t buffer_bytes, uint64_t lines_to_read, double* gbytes_per_s);
/* the same for random aligned loads of granule_bytes (32, 64 or 128) each */
int rs_measure_random_access_granule(int device, uint64_t buffer_bytes, uint64_t lines_to_read, int granule_bytes,
                                     double* gbytes_per_s);
const char* rs_last_error(void);
/* pinned host memory for the host-pointer calls (optional; any host memory is accepted) */
int rs_host_alloc(void** p, size_t bytes);
void rs_host_free(void* p);

/* ---- class BWT (include/bwt/bwt.h:6-15), batched ------------------------------------------------- */
uint64_t rs_get_pc(const rs_index* ix, char b);   /* RLEBWT::getPC,    rlebwt.cpp:229-231 */
uint64_t rs_get_bwlen(const rs_index* ix);        /* RLEBWT::getBWLen, rlebwt.cpp:303-305 */
/* RLEBWT::getOcc(b, index), rlebwt.cpp:268-301: out[i] = occurrences of b[i] in BWT[0..index[i]] */
int rs_get_occ_batch(const rs_index* ix, const char* b, const uint64_t* index, size_t n, uint64_t* out);
/* RLEBWT::getChar(index), rlebwt.cpp:202-227 */
int rs_get_char_batch(const rs_index* ix, const uint64_t* index, size_t n, char* out);
/* RLEBWT::getF(index), rlebwt.cpp:307-314 */
int rs_get_f_batch(const rs_index* ix, const uint64_t* index, size_t n, char* out);

/* RLEBWT::getOccAt(b, count), rlebwt.cpp:233-266: position of the count-th occurrence of b
 * (count outside [1, occurrences of b] yields UINT64_MAX; undefined in the reference) */
int rs_get_occ_at_batch(const rs_index* ix, const char* b, const uint64_t* count, size_t n, uint64_t* out);

/* ---- include/bwt/query.h:18-32, batched ---------------------------------------------------------- */
/* findInterval (src/bwt/query.cpp:24-41) for Q k-mers of one length k stored back to back.
 * lower/upper are the reference's BWTInterval, including the value at the step where the search
 * stopped early (query.cpp:35-37). */
int rs_find_interval_batch(const rs_index* ix, const char* kmers, int k, size_t Q, uint64_t* lower, uint64_t* upper);
/* findInterval for Q strings of any length: string q = text[offsets[q] .. offsets[q+1]).
 * An empty string (undefined behaviour in the reference) yields lower = 1, upper = 0. */
int rs_find_interval_var_batch(const rs_index* ix, const char* text, const uint64_t* offsets, size_t Q,
                               uint64_t* lower, uint64_t* upper);
int rs_find_interval_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, void* d_lower,
                               void* d_upper, void* stream);
int rs_find_interval_var_batch_dev(const rs_index* ix, const void* d_text, const void* d_offsets, size_t Q,
                                   void* d_lower, void* d_upper, void* stream);

/* count_reads (src/service/service.cpp:279-315): per k-mer, the number of BWT rows matching it on the
 * forward strand (strand = 0) or its reverse complement (strand = 1); a k-mer with any byte outside
 * ACGT counts 0 (service.cpp:299-301). */
int rs_count_batch(const rs_index* ix, const char* kmers, int k, size_t Q, int strand, uint64_t* counts);
int rs_count_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, int strand, void* d_counts,
                       void* stream);

/* rs_count_batch_dev for callers that may not allocate device memory (code under CUDA stream capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arguments, same answers as rs_count_batch_dev; the call itself performs no cudaMalloc / cudaFree /
 * allocating library routine: its scratch (8 bytes of packed key per query + the work counter of the search kernel)
 * is a file-scope `static __device__` array of the library, RS_NOALLOC_MAX_QUERIES * 8 + 256 bytes (8 MiB) that every
 * GPU loading the library holds until the process exits.  Limits: Q <= RS_NOALLOC_MAX_QUERIES and 1 <= k <= 31
 * (fixed-length k-mers); beyond them the call returns RS_STATUS_ARG with a message and launches nothing.  The index's
 * device must be the current device.  All calls on one GPU share the array: calls that overlap on the GPU (different
 * streams with no ordering between them) RACE; calls on one stream, or ordered by events, are safe. */
#define RS_NOALLOC_MAX_QUERIES (1u << 20)
int rs_count_batch_noalloc_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, int strand, void* d_counts,
                               void* stream);

/* ---- locate -------------------------------------------------------------------------------------- */
/* Attach the lexicographic-rank -> read-index table (SGA's .sai, written by `sga index` next to the
 * .bwt: submodules/sga/src/SuffixTools/SampledSuffixArray.cpp:171-189).  sai has num_strings entries. */
int rs_index_set_sai(rs_index* ix, const uint64_t* sai, uint64_t num_strings);
int rs_index_set_sai_dev(rs_index* ix, const void* d_sai, uint64_t num_strings); /* device-to-device copy */
/* extractPrefix (src/bwt/query.cpp:43-63) for n BWT rows: LF-walk to the '$'.
 *   lexrank[i]  terminal row = lexicographic rank of the read (UINT64_MAX if the walk did not end
 *               within max_steps bases: inside the reference's tail region it can cycle forever)
 *   steps[i]    bases walked (UINT32_MAX on cut-off)            -- may be NULL
 *   read_index[i]  sai[lexrank[i]]                               -- may be NULL; needs rs_index_set_sai
 *   bases       n x max_steps chars, the extracted prefix in reading order, NUL padded -- may be NULL */
int rs_extract_prefix_batch(const rs_index* ix, const uint64_t* rows, size_t n, uint32_t max_steps,
                            uint64_t* lexrank, uint32_t* steps, uint64_t* read_index, char* bases);
int rs_extract_prefix_batch_dev(const rs_index* ix, const void* d_rows, size_t n, uint32_t max_steps,
                                void* d_lexrank, void* d_steps, void* d_read_index, void* d_bases, void* stream);

/* extractPostfix (src/bwt/query.cpp:65-85) for n BWT rows: forward walk by getF / getOccAt to the '$'.
 *   read_index[i] the row (< num_strings) at which getF returned '$' = the read's index in builder
 *                 input order (UINT64_MAX if the walk did not end within max_steps bases)
 *   steps[i]      bases walked (UINT32_MAX on cut-off)                -- may be NULL
 *   bases         n x max_steps chars in reading order, NUL padded    -- may be NULL */
int rs_extract_postfix_batch(const rs_index* ix, const uint64_t* rows, size_t n, uint32_t max_steps,
                             uint64_t* read_index, uint32_t* steps, char* bases);
int rs_extract_postfix_batch_dev(const rs_index* ix, const void* d_rows, size_t n, uint32_t max_steps,
                                 void* d_read_index, void* d_steps, void* d_bases, void* stream);

/* ---- count + locate: the per-row loop of query() (src/bwt/query.cpp:87-100) without the strings ---- */
/* For Q k-mers: findInterval, then for every row of each interval (at most max_hits rows per k-mer)
 * extractPrefix's walk to the '$' and the .sai lookup.
 *   hit_offsets[Q+1]   CSR offsets of each k-mer's hits (always written)
 *   capacity           entries available in rows / lexrank / read_index / steps; if the hits do not
 *                      fit, only hit_offsets is written and RS_STATUS_CAPACITY is returned
 *   rows, lexrank, read_index, steps   per hit, any of them may be NULL (read_index needs the .sai) */
int rs_locate_batch(const rs_index* ix, const char* kmers, int k, size_t Q, uint64_t max_hits, uint32_t max_steps,
                    uint64_t* hit_offsets, uint64_t capacity, uint64_t* rows, uint64_t* lexrank,
                    uint64_t* read_index, uint32_t* steps);
/* Same with device buffers; *total_hits (host) receives hit_offsets[Q]. Synchronises `stream` once. */
int rs_locate_batch_dev(const rs_index* ix, const void* d_kmers, int k, size_t Q, uint64_t max_hits,
                        uint32_t max_steps, void* d_hit_offsets, uint64_t capacity, void* d_rows, void* d_lexrank,
                        void* d_read_index, void* d_steps, uint64_t* total_hits, void* stream);

/* ---- sample sets --------------------------------------------------------------------------------- */
/* Attach the read-index -> sample-codes table that stands in for the 64 RocksDB stores on this path
 * (key = read string, value = concatenated codes: src/service/service.cpp:816-845; the table is the
 * same values addressed by read index instead of read string).  Read r owns codes
 * [offsets[r], offsets[r+1]) of size_of_sample bytes each; offsets has num_strings+1 entries. */
int rs_index_set_samples(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, uint64_t num_strings,
                         int size_of_sample);
int rs_index_set_samples_dev(rs_index* ix, const void* d_offsets, const void* d_codes, uint64_t num_strings,
                             int size_of_sample); /* device-to-device copy */
/* With the store's other metadata (has_other_meta_data = 1, demo/TEMPLATE.service.cfg:23): meta holds 2 bytes per
 * code, the count and lane characters that follow it in the value (service.cpp:829-843); NULL = none. */
int rs_index_set_samples_meta(rs_index* ix, const uint64_t* offsets, const uint8_t* codes, const uint8_t* meta,
                              uint64_t num_strings, int size_of_sample);
/* The same table straight from the ETL's merged read file (`read\n codes\n` pairs in RLO order, the input of
 * src/util/load_data_into_rocksdb.cpp:41-49): the i-th read without 'N' is read index i (build_bwt.sh:919);
 * with has_other_meta_data each code is followed by count and lane bytes (service.cpp:829-843), kept beside the
 * codes (rs_samples_meta_batch).  Replaces the 64 RocksDB loads of build_bwt.sh step 9 on this path. */
int rs_index_load_samples_file(rs_index* ix, const char* path, int size_of_sample, int has_other_meta_data);
/* Sample codes of n reads, order and duplicates preserved (service.cpp:829-843):
 * out_offsets[n+1] in codes (always written); out_codes holds `capacity` codes, else RS_STATUS_CAPACITY. */
int rs_samples_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint64_t* out_offsets,
                     uint64_t capacity, uint8_t* out_codes);
/* The same plus, per code, its count and lane bytes as stored (char - 33 each is the number: service.cpp:833-841);
 * out_meta = 2 x capacity bytes, needs a table attached with metadata. */
int rs_samples_meta_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint64_t* out_offsets,
                          uint64_t capacity, uint8_t* out_codes, uint8_t* out_meta);
/* Carrier sets: for each of Q k-mers whose hits are read_index[hit_offsets[q] .. hit_offsets[q+1]),
 * the distinct sample codes over those reads in ascending order. */
int rs_carriers_batch(const rs_index* ix, const uint64_t* hit_offsets, const uint64_t* read_index, size_t Q,
                      uint64_t* out_offsets, uint64_t capacity, uint8_t* out_codes);

/* ---- read strings (output=reads) ------------------------------------------------------------------
 * The reference rebuilds each matching read by walking the BWT from the hit row (query(): extractPrefix +
 * extractPostfix, src/bwt/query.cpp:87-100; find_reads, service.cpp:714-751).  With the read index from the
 * locate stage the string can be read from a table: all reads in builder input order, 2 bits per base.
 *   rs_index_set_reads        text = the reads concatenated (ACGT only), offsets[num_strings + 1]
 *   rs_index_load_reads_file  the same from the ETL's merged `read / codes` file (reads with 'N' skipped)
 *   rs_reads_batch            out = n x max_len bytes, zero padded; out_len[i] = the read's length
 *                             (2^32-1 for a read index outside the table)                              */
int rs_index_set_reads(rs_index* ix, const char* text, const uint64_t* offsets, uint64_t num_strings);
int rs_index_load_reads_file(rs_index* ix, const char* path);
int rs_reads_batch(const rs_index* ix, const uint64_t* read_index, size_t n, uint32_t max_len, char* out, uint32_t* out_len);
int rs_reads_batch_dev(const rs_index* ix, const void* d_read_index, size_t n, uint32_t max_len, void* d_out,
                       void* d_out_len, void* stream);

/* ---- k-mer tiling of long queries (/kmermatch) ---------------------------------------------------
 * get_tiles + the findInterval of find_reads for a whole batch (src/service/service.cpp:232-246, 466-500):
 * query q (text[offsets[q] .. offsets[q+1])) is cut into tiles of `kmer` bases starting at 0, skip+1,
 * 2(skip+1), ...; tile_offsets (Q+1, out) is the CSR of tiles per query; lower/upper (capacity entries)
 * receive one interval per tile.  Tiles holding a byte outside ACGT, which the reference skips, come back
 * as (1, 0).  RS_STATUS_CAPACITY when there are more tiles than `capacity` (tile_offsets is still filled). */
int rs_tile_intervals_batch(const rs_index* ix, const char* text, const uint64_t* offsets, size_t Q, uint32_t kmer,
                            uint32_t skip, uint64_t* tile_offsets, uint64_t capacity, uint64_t* lower, uint64_t* upper);

/* ---- /gt seed stage: find_gt_reads (src/service/service.cpp:507-711) for a batch of site queries ---- */
/* text/offsets: Q windows, already oriented and cut around the site as GtTask::run does (service.cpp:1027-1037);
 * pos[q]: the site, 1-based, in window q.  Every window is cut into tiles of `kmer` bases every skip+1 bases
 * ((len - kmer) / (skip + 1) + 1 of them, :1049); per tile the reference's interval search with its extension loop
 * (while the interval holds more than max_interval_size rows: the constant 10000 of service.cpp:91) and its reach
 * test on the rows' prefix / postfix lengths (indel allowance 4).  The union over a window's tiles comes back as
 * sorted unique read indices (builder input order, through the .sai):
 *   out_offsets[Q+1]  CSR offsets, always written when the call gets that far
 *   read_index        capacity entries; RS_STATUS_CAPACITY if the sets do not fit
 *   flags[Q]          RS_GT_FLAG_*: tiles on which the reference itself throws (THROW) or never returns (HANG)
 *                     contribute no reads and mark their query; WALK marks rows whose walk was cut at max_steps
 *   max_rows          bound on the candidate rows (before the reach test) the call may expand on the device;
 *                     *candidate_rows (may be NULL) receives the number needed
 * Needs the .sai table; postfix lengths come from the packed read table when one is attached (and rank is exact),
 * else from the select walk. */
#define RS_GT_KEEP_ALL 0u
#define RS_GT_NEED_POSTFIX 1u
#define RS_GT_NEED_PREFIX 2u
#define RS_GT_FLAG_THROW 1u
#define RS_GT_FLAG_HANG 2u
#define RS_GT_FLAG_WALK 4u
int rs_gt_reads_batch(const rs_index* ix, const char* text, const uint64_t* offsets, const uint64_t* pos, size_t Q,
                      uint32_t kmer, uint32_t skip, uint64_t max_interval_size, uint32_t max_steps, uint64_t max_rows,
                      uint64_t* out_offsets, uint64_t capacity, uint64_t* read_index, uint32_t* flags,
                      uint64_t* candidate_rows);

/* ---- population queries over a set of shards on one GPU ------------------------------------------------
 * The reference runs one worker per BWT shard; its front end broadcasts every query to all of them and, per query,
 * sums their counts and concatenates their read lists (src/service/server.cpp:124, 184-197, 404-410, 957); a worker
 * answers with findInterval, one walk per hit row and one sample-store lookup per read (service.cpp:279-315, 714-751,
 * 816-845).  A shard set holds the shards resident on one GPU and answers a whole batch in one device pipeline.
 *
 *   rs_shardset_add     the index stays owned by the caller and must outlive the set; needs its .sai attached
 *                       (and its sample table, if carrier sets are wanted); shard_id < 4096 names it in the hits
 *   rs_shardset_seal    locate_every = 1: builds a read-index table per shard (4 bytes per BWT row: the rows of an
 *                       interval are contiguous, so a query's read indices are read with coalesced loads instead of
 *                       walking; per-read sample-set bitmaps as well when the tables hold <= 128 samples);
 *                       locate_every > 1: the sampled locate table of rs_index_build_locate_samples per shard;
 *                       locate_every = 0: plain walks (K2).  max_read_len bounds every walk.
 *   rs_pop_query[_dev]  Q k-mers of one length k <= 31: ASCII (packed = 0, k bytes each) or packed (packed = 1:
 *                       one uint64 per k-mer, 2 bits per base A,C,G,T = 0..3, first base most significant, bit 63
 *                       set = the k-mer holds a byte outside ACGT and counts 0, service.cpp:299-301).  strand = 0 searches
 *                       the k-mers as given, 1 their reverse complements (the worker's two tasks per request).
 *     counts[Q]           population count = sum over the set's shards of the interval widths (count_reads)
 *     hit_offsets[Q+1]    may be NULL (count only).  CSR of each k-mer's hits; a k-mer's hits are contiguous: shard
 *                         after shard in the order added, rows of the shard's interval in order, at most max_hits
 *                         rows per shard
 *     hits[capacity]      shard_id << 40 | read index (the read's position in the shard's builder input order, what
 *                         the .sai gives for the terminal row of extractPrefix's walk); UINT64_MAX for a row whose
 *                         walk does not end (reference tail region).  Hits beyond capacity are dropped:
 *                         *total_hits > capacity tells (the host-pointer call then returns RS_STATUS_CAPACITY)
 *     carriers[Q x W]     may be NULL.  Bit s of a k-mer's W 32-bit words = sample number s (0-based: code - 1,
 *                         create_samples_hash.cpp:41-50) carries at least one of its hit reads; W = carrier_words
 *                         of rs_shardset_get_info.  Needs the hit list.
 * The _dev call is asynchronous on `stream` and never synchronises; calls on one set are serialised.  The host-pointer
 * calls copy the k-mers in, run the pipeline and copy the answers out -- range by range of the batch while the locate
 * stage works on the next range (large batches are sorted by (range, k-mer) for that) -- and return when the answers
 * are in the caller's buffers; pinned buffers (rs_host_alloc) make the copies asynchronous. */
typedef struct rs_shardset rs_shardset;
typedef struct rs_shardset_info {
  uint32_t num_shards, sealed, locate_every, max_read_len, num_samples, carrier_words;
  uint64_t index_bytes;   /* blocks, directories and attached tables of the shards            */
  uint64_t table_bytes;   /* read-index tables and sample-set bitmaps built by the seal       */
  uint64_t bwlen, num_strings, num_runs; /* summed over the shards                            */
  uint64_t chunk_queries; /* queries per pipeline chunk                                        */
} rs_shardset_info;
int rs_shardset_create(int device, rs_shardset** out);
int rs_shardset_add(rs_shardset* set, rs_index* ix, uint32_t shard_id);
int rs_shardset_seal(rs_shardset* set, uint32_t max_read_len, uint32_t locate_every);
void rs_shardset_destroy(rs_shardset* set);
int rs_shardset_get_info(const rs_shardset* set, rs_shardset_info* info);
int rs_shardset_enable_stats(rs_shardset* set, int on);
/* work counters of the set's calls: executed updateInterval steps (2 getOcc each), LF steps of the walks, and the
 * 64-byte blocks K1 read (a step whose two positions share a block reads one) */
int rs_shardset_read_stats(rs_shardset* set, uint64_t* interval_steps, uint64_t* walk_steps, uint64_t* block_loads,
                           int reset);
/* Measurement aid: CUDA events around the pipeline's stages on the call's own stream.  ms[4] / launches[4] =
 * time and kernel launches of 0 k_pack_kmers, 1 K1 (search), 2 offsets, 3 locate + sample sets, summed over the
 * calls since the last read (which synchronises the device). */
int rs_shardset_enable_timing(rs_shardset* set, int on);
int rs_shardset_read_timing(rs_shardset* set, double ms[4], uint64_t launches[4]);
int rs_pop_query_dev(rs_shardset* set, const void* d_kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                     void* d_counts, void* d_hit_offsets, void* d_hits, uint64_t capacity, void* d_carriers,
                     void* d_total_hits, void* stream);
int rs_pop_query(rs_shardset* set, const void* kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                 uint64_t* counts, uint64_t* hit_offsets, uint64_t* hits, uint64_t capacity, uint32_t* carriers,
                 uint64_t* total_hits);
/* rs_pop_query with the hit list in 5 bytes per hit instead of 8 (the list is most of what the call copies to the
 * host: 790 M hits per 10 M 31-mers on BASELINE configs[2]): hit i = the 40-bit little-endian value at hits40 + 5 * i,
 * shard id << 32 | read index; 2^40 - 1 where rs_pop_query reports 2^64 - 1.  hits40 holds 5 * capacity bytes.
 * RS_STATUS_TOO_LARGE when a shard id >= 255 or a read index >= 2^32 occurs (the answers other than the list are valid). */
int rs_pop_query_hits40(rs_shardset* set, const void* kmers, int k, size_t Q, int packed, int strand, uint64_t max_hits,
                        uint64_t* counts, uint64_t* hit_offsets, uint8_t* hits40, uint64_t capacity, uint32_t* carriers,
                        uint64_t* total_hits);

/* ---- population queries across the GPUs of one box ---------------------------------------------------------
 * One process per GPU, each with a shard set over the shards it holds.  Every rank runs the whole batch on its set;
 * the answers are fanned in -- by the pipeline's own kernels, through NVLink peer memory (CUDA IPC), no collective
 * library -- to the rank that OWNS the k-mer (its first four bases as a base-4 number modulo the ranks):
 *   counts     each rank adds its shards' count into the owner's (the front end's sum, server.cpp:184-197, 404-410)
 *   hit lists  each rank reports its number of hits per k-mer, the owner lays the list out (rank 0's hits first,
 *              then rank 1's, ...) and tells every rank its slot; the locate kernel then writes the read indices
 *              straight into the owner's list (the front end's concatenation of read lists)
 *   carriers   each rank ORs the sample set of its hit reads into the owner's bitmap
 * rs_popcomm_create allocates this rank's result buffers (max_queries per call, hit_capacity hits);
 * export / connect exchange the 64-byte IPC handles as for rs_comm.  rs_pop_query_comm_dev is asynchronous on
 * `stream`; all ranks must call it with the same batch and arguments.  The results stay in the communicator's
 * buffers until its next call: rs_popcomm_results gives their device addresses (layouts as rs_pop_query_dev;
 * entries of k-mers owned by other ranks are 0 / empty).  Carrier sets across GPUs need <= 128 samples. */
typedef struct rs_popcomm rs_popcomm;
int rs_popcomm_create(int device, int rank, int world, uint64_t max_queries, uint64_t hit_capacity, rs_popcomm** out);
int rs_popcomm_export(rs_popcomm* c, uint8_t handle[64]);
int rs_popcomm_connect(rs_popcomm* c, const uint8_t* handles);
void rs_popcomm_destroy(rs_popcomm* c);
int rs_pop_query_comm_dev(rs_shardset* set, rs_popcomm* comm, const void* d_kmers, int k, size_t Q, int packed, int strand,
                          uint64_t max_hits, int want_hits, int want_carriers, void* stream);
int rs_popcomm_results(rs_popcomm* c, void** d_counts, void** d_hit_offsets, void** d_hits, void** d_carriers,
                       void** d_total_hits, uint64_t* hit_capacity);
int rs_popcomm_error(rs_popcomm* c, int* timed_out);

/* ---- fan-in of per-shard counts across the GPUs of one box ----------------------------------------
 * Replaces the front end's sum over shard replies (src/service/server.cpp:184-197, 404-410).  One
 * process per GPU; every rank searches the whole batch in its shards and each finished query's count
 * is added by the search kernel itself into the buffer of the rank that OWNS the k-mer (its first four
 * bases as a base-4 number, A,C,G,T = 0..3, modulo the number of ranks) through NVLink peer memory.
 *
 *   rs_comm_create   allocates this rank's buffers (capacity max_queries per step)
 *   rs_comm_export   64-byte CUDA IPC handle of that allocation, to be exchanged by the caller
 *   rs_comm_connect  maps the peers: `handles` = world x 64 bytes in rank order (own slot ignored)
 *   per step, identically on every rank and on one stream:
 *     rs_comm_begin(Q) ; rs_count_accumulate_dev(...) once per local shard ; rs_comm_finish
 *   rs_comm_finish   cross-GPU barrier on the stream; *d_counts = device array of Q u64 whose entries for
 *                    the k-mers this rank owns hold the population counts (other entries are 0); valid
 *                    until the next-but-one rs_comm_begin (three buffers rotate), for readers ordered after the
 *                    barrier on `stream` (a reader on another stream must wait for an event recorded there)
 *   rs_comm_error    1 if a barrier gave up waiting for a peer (5 s)
 * Fixed-length queries of 1..31 characters; any byte outside ACGT counts 0 (service.cpp:299-301). */
typedef struct rs_comm rs_comm;
#define RS_COMM_HANDLE_BYTES 64
int rs_comm_create(int device, int rank, int world, uint64_t max_queries, rs_comm** out);
int rs_comm_export(rs_comm* c, uint8_t handle[RS_COMM_HANDLE_BYTES]);
int rs_comm_connect(rs_comm* c, const uint8_t* handles);
void rs_comm_destroy(rs_comm* c);
int rs_comm_begin(rs_comm* c, uint64_t Q, void* stream);
int rs_count_accumulate_dev(const rs_index* ix, rs_comm* c, const void* d_kmers, int k, size_t Q, int strand,
                            void* stream);
int rs_comm_finish(rs_comm* c, void* stream, void** d_counts);
int rs_comm_error(rs_comm* c, int* timed_out);

#ifdef __cplusplus
}
#endif
#endif
