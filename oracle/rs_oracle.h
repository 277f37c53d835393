/* TEST INFRASTRUCTURE ONLY -- the CPU oracle ("port") for the ReadServer FM-index hot path.
 *
 * Plain-C restatement of the reference algorithm (wtsi-svi/ReadServer, include/bwt/{BPTree,BPNodes,
 * rlunit,alphabet}.h, src/bwt/{rlebwt,rlebwt_reader,query}.cpp).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this; the product path never does.
 *
 * Parity of this port is PINNED: tests/test_oracle_vs_ref.py checks it call-for-call against the
 * unmodified reference compiled into oracle/_ref/libref_bwt.so (getOcc/getChar/getOccAt/getF at
 * every position of small BWTs, findInterval, extractPrefix/extractPostfix, and a byte-for-byte
 * comparison of the .bpi2 file with the one written by the reference's index_rlebwt), and
 * tests/golden/ holds vectors generated from the reference itself.
 */
#ifndef RS_ORACLE_H
#define RS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rso_index rso_index;

/* rlebwt_reader.cpp:27-48 + rlebwt.cpp:11-32.  Returns 0, or 1 (cannot open), 2 (bad magic),
 * 3 (short file), 4 (run byte with symbol code > 4 or zero length). If use_bpi2 != 0 and
 * <path>.bpi2 exists the tree is read from it (rlebwt.cpp:163-200) instead of being built. */
int rso_open(const char* bwt_path, int use_bpi2, rso_index** out);
int rso_from_runs(const uint8_t* runs, uint64_t nruns, uint64_t num_strings, uint64_t num_symbols,
                  rso_index** out);
void rso_close(rso_index* ix);
/* rlebwt.cpp:150-161 + BPTree.h:189-199 + BPNodes.h:198-214 */
int rso_write_bpi2(const rso_index* ix, const char* path);

uint64_t rso_bwlen(const rso_index* ix);
uint64_t rso_num_strings(const rso_index* ix);
uint64_t rso_num_runs(const rso_index* ix);
uint64_t rso_depth(const rso_index* ix);
uint64_t rso_vsum_size(const rso_index* ix);
/* number of times a lookup would have read m_vSum out of bounds; must stay 0 */
uint64_t rso_q2_violations(const rso_index* ix);

uint64_t rso_get_pc(const rso_index* ix, char b);                    /* rlebwt.cpp:229-231 */
uint64_t rso_get_occ(const rso_index* ix, char b, uint64_t index);   /* rlebwt.cpp:268-301 */
uint64_t rso_get_occ_at(const rso_index* ix, char b, uint64_t bc);   /* rlebwt.cpp:233-266 */
char rso_get_char(const rso_index* ix, uint64_t index);              /* rlebwt.cpp:202-227 */
char rso_get_f(const rso_index* ix, uint64_t index);                 /* rlebwt.cpp:307-314 */

/* query.cpp:24-41 (w must be non-empty: SURVEY.md Q3) */
void rso_find_interval(const rso_index* ix, const char* w, size_t len, uint64_t* lower, uint64_t* upper);
/* query.cpp:43-63: writes the prefix (already reversed into reading order) into out (cap bytes incl.
 * NUL), returns its length; *terminal = idx after the '$' step = lexicographic rank of the read. */
size_t rso_extract_prefix(const rso_index* ix, uint64_t row, char* out, size_t cap, uint64_t* terminal,
                          size_t max_steps);
/* query.cpp:65-85: *terminal = the row (< num_strings) at which getF returned '$' = the read's index in
 * builder input order ($_i < $_j <=> i < j).  The reference also performs the select of that step
 * (query.cpp:73-74) before it breaks; so does this port, but the row reported is the one before it. */
size_t rso_extract_postfix(const rso_index* ix, uint64_t row, char* out, size_t cap, uint64_t* terminal,
                           size_t max_steps);
/* NOTE: a walk that enters the reference's tail-clamp region can cycle forever; the
 * reference itself would hang there.  Walkers stop after max_steps bases, return (size_t)-1 and set
 * *terminal = UINT64_MAX (batch form: steps = UINT32_MAX). */

/* Batch drivers (threads stride the array, as BASELINE.md section 3 describes). Return seconds. */
double rso_get_occ_batch(const rso_index* ix, const char* b, const uint64_t* idx, size_t n, uint64_t* out);
double rso_get_char_batch(const rso_index* ix, const uint64_t* idx, size_t n, char* out);
double rso_get_occ_at_batch(const rso_index* ix, const char* b, const uint64_t* cnt, size_t n, uint64_t* out);
double rso_find_interval_batch(const rso_index* ix, const char* kmers, int k, size_t Q, uint64_t* lower,
                               uint64_t* upper, int nthreads);
double rso_locate_batch(const rso_index* ix, const uint64_t* rows, size_t n, uint64_t* lexrank,
                        uint32_t* pre_steps, uint64_t* read_index, uint32_t* post_steps, int nthreads,
                        uint32_t max_steps);

/* Sample value decode, service.cpp:816-845: value = repeated { code[size_of_sample] [, count+33, lane+33] }.
 * Writes up to cap records; returns the number of records in the value. codes_out gets
 * size_of_sample bytes per record, counts/lanes (may be NULL) get the decoded metadata. */
size_t rso_decode_sample_value(const uint8_t* value, size_t value_len, int size_of_sample,
                               int has_other_meta_data, uint8_t* codes_out, int32_t* counts,
                               int32_t* lanes, size_t cap);
/* create_samples_hash.cpp:31-51: index -> k base-223 digits offset by 33 (most significant first). */
void rso_sample_code(uint64_t sample_idx, int size_of_sample, uint8_t* out);

#ifdef __cplusplus
}
#endif
#endif
