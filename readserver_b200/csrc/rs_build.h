/* Host-side re-encoder: SGA run bytes (.bwt) -> 64-byte blocks + directories (rs_layout.h).
 * Takes the place of RLEBWT::initialiseFMIndex (src/bwt/rlebwt.cpp:34-148) and of the .bwt reader
 * (src/bwt/rlebwt_reader.cpp:21-48).  One streaming pass over the run bytes.
 */
#ifndef RS_BUILD_H
#define RS_BUILD_H

#include <stdint.h>
#include <string>
#include <vector>

#include "rs_layout.h"

struct RsHostIndex {
  std::vector<rs_block> blocks;
  std::vector<uint64_t> dir;
  std::vector<uint32_t> sel[RS_NSYM];
  uint32_t sel_shift[RS_NSYM] = {0, 0, 0, 0, 0};
  uint64_t n = 0, num_strings = 0, num_runs = 0;
  uint64_t pc[RS_NSYM] = {0, 0, 0, 0, 0};
  uint64_t tot[RS_NSYM] = {0, 0, 0, 0, 0};
  uint64_t init_occ[RS_NSYM] = {0, 0, 0, 0, 0};
  uint64_t clamp_c = 0, clamp_char = 0, vlast = 0;
  uint64_t depth = 0;  /* depth the reference's BP-tree would have (rlebwt.cpp:51-72) */
  bool q2_safe = true; /* false: the reference would read m_vSum out of bounds */
};

enum {
  RS_OK = 0,
  RS_ERR_OPEN = 1,      /* cannot open file                                       */
  RS_ERR_MAGIC = 2,     /* not an SGA binary RLE BWT (rlebwt_reader.cpp:31-34)    */
  RS_ERR_SHORT = 3,     /* truncated file                                         */
  RS_ERR_RUNBYTE = 4,   /* run byte with symbol > 4 or length 0                   */
  RS_ERR_ARG = 5,       /* null / empty / inconsistent argument                   */
  RS_ERR_CUDA = 6,      /* CUDA runtime error (message via rs_last_error)         */
  RS_ERR_TOO_LARGE = 7, /* more than 2^32 blocks in one shard                     */
  RS_ERR_NO_TABLE = 8   /* locate/sample call without the table it needs          */
};

/* Reads header + run bytes of an SGA .bwt. */
int rs_read_bwt_file(const char* path, std::vector<uint8_t>& runs, uint64_t& num_strings, uint64_t& num_symbols);

/* Builds blocks, directories and constants from run bytes. */
int rs_build_host_index(const uint8_t* runs, uint64_t nruns, uint64_t num_strings, uint64_t num_symbols,
                        RsHostIndex& out);

/* Sample table from the ETL's merged read file (`read\n codes\n` pairs in RLO order: rlosort_seq_and_convert_
 * sample_names.cpp:47-53, merge_rlosorted_reads.cpp:45-47; the file load_data_into_rocksdb.cpp:41-49 reads).
 * Reads holding 'N' never reach the BWT (build_bwt.sh:919), so the i-th N-free read of the file is read
 * index i.  Each value is a run of size_of_sample-byte codes, each followed by 2 bytes (count, lane) iff
 * has_other_meta_data (service.cpp:829-843); the codes go to `codes`, the 2 metadata bytes per code to
 * `meta_out` when given.  offsets gets one entry per kept
 * read + 1.  Returns RS_OK, RS_ERR_OPEN, or RS_ERR_ARG (odd line count / value length not a multiple). */
int rs_parse_etl_samples(const char* path, int size_of_sample, int has_other_meta_data, std::vector<uint64_t>& offsets,
                         std::vector<uint8_t>& codes, std::vector<uint8_t>* meta_out = nullptr);

/* The reads of the same file that reach the BWT (odd lines without 'N'), concatenated, with their offsets. */
int rs_parse_etl_reads(const char* path, std::string& text, std::vector<uint64_t>& offsets);

/* Scalar fields of the kernel-side view from a host index (the caller sets the pointers), and the
 * reference-tail switch: folds "last m_vSum window and at/after the clamp" into one threshold per
 * lookup kind, and picks initInterval's constant (query.cpp:18-21). */
void rs_fill_view(const RsHostIndex& h, rs_index_dev* d);
void rs_view_set_tail(const RsHostIndex& h, bool on, rs_index_dev* d);
void rs_view_set_tail(const uint64_t init_occ_ref[RS_NSYM], bool on, rs_index_dev* d);

#endif
