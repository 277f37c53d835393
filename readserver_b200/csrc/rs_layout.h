/* HBM layout of one population-BWT shard (shared by the host builder and the sm_100a kernels).
 *
 * Replaces the reference's RLVector + multi-level BP-tree (include/bwt/rlunit.h:8-89,
 * include/bwt/BPTree.h:20-206, BPNodes.h:34-241, built by src/bwt/rlebwt.cpp:34-148) with a flat
 * array of 64-byte blocks -- two 32-byte DRAM sectors, half an L2 line, read with four independent
 * 128-bit loads -- and a position directory that resolves "symbol position -> block" in one 8-byte
 * read.  A rank is two dependent memory accesses: directory entry, then block.
 *
 *   block (64 B, 64-B aligned)
 *     +0   u64 start          symbols before this block
 *     +8   u32 cnt_lo[4]      low 32 bits of the occurrences of A,C,G,T before this block
 *     +24  u32 cnt_hi         byte c-1 = bits 32..39 of that count ('$' = start - sum)
 *     +28  u8  runs[36]       SGA run bytes  sym<<5 | len  (rlunit.h:8-16), 0x00 = padding
 *
 *   position directory: one u64 per 2^RS_SLOT_SHIFT symbols
 *     bits  0..31  index of the block holding the slot's first symbol
 *     bits 32..40  b1  offset inside the slot at which block +1 starts (0 = no such boundary)
 *     bits 41..49  b2  ... block +2
 *     bits 50..58  b3  ... block +3
 *     bit  62      more than three block boundaries fall inside the slot (then the start fields of
 *                  the following blocks are compared until the block is found)
 *
 *   select directory (per symbol c): one u32 per 2^s occurrences of c, s = sel_shift[c] chosen so
 *     that a directory slot spans about two blocks (a sparse symbol such as '$' gets a small s)
 *     entry k = index of the block holding the ((k << s) + 1)-th occurrence of c,
 *     i.e. the block with cnt_before[c] <= k << s < cnt_before_next[c].
 *
 * Limits of one shard: 2^32 blocks (256 GB), 2^40 occurrences per symbol.
 */
#ifndef RS_LAYOUT_H
#define RS_LAYOUT_H

#include <stdint.h>

#define RS_BLOCK_BYTES 64
#define RS_HDR_BYTES 28
#define RS_PAYLOAD 36 /* runs per block */
#define RS_BLOCK_WORDS (RS_PAYLOAD / 4)
#ifndef RS_SLOT_SHIFT
#define RS_SLOT_SHIFT 9 /* 512-symbol directory slots */
#endif
#define RS_SLOT_SIZE (1u << RS_SLOT_SHIFT)
#define RS_SLOT_MASK (RS_SLOT_SIZE - 1u)
#define RS_DIR_MORE (1ull << 62)
#define RS_SEL_SHIFT_MAX 10
#define RS_NSYM 5
#define RS_NO_ROW 0xFFFFFFFFFFFFFFFFull

typedef struct rs_block {
  uint64_t start;
  uint32_t cnt_lo[4];
  uint32_t cnt_hi;
  uint8_t runs[RS_PAYLOAD];
} rs_block;

/* Everything a kernel needs, passed by value as a __grid_constant__ parameter. */
typedef struct rs_index_dev {
  const void* blocks;        /* rs_block[nblocks]                                              */
  const uint64_t* dir;       /* position directory, ndir entries                                */
  const uint32_t* sel[RS_NSYM]; /* select directories                                           */
  uint64_t n;                /* BWT length in symbols (getBWLen)                                */
  uint64_t nblocks;
  uint64_t num_strings;
  uint64_t pc[RS_NSYM];      /* C[] array: getPC (rlebwt.cpp:229-231)                           */
  uint64_t tot[RS_NSYM];     /* occurrences of each symbol in the whole BWT                     */
  uint64_t init_occ[RS_NSYM];/* what the reference's getOcc(c, n-1) returns (initInterval)      */
  /* Reference tail behaviour (BPTree.h:74-94, 135-157): for positions i with
   * ((i+1) >> 16) >= vlast and i >= clamp the reference answers as if i were clamp - 1.
   * rank (getOcc) is confined to the top-level bucket of the last m_vSum hint; access (getChar)
   * to the first bucket on that hint's path whose index is non-zero (BPTree.h:148-151 tests
   * `offset == 0`), which is a deeper, smaller bucket when the top-level index is 0. */
  uint64_t clamp_c;          /* getOcc  */
  uint64_t clamp_char;       /* getChar */
  uint64_t vlast;
  /* the same rule folded into one compare per lookup: i >= tail_from  <=>  the clamp applies
   * (max(vlast * 65536 - 1, clamp) when emulating the reference, 2^64-1 = never otherwise) */
  uint64_t tail_from;        /* getOcc  */
  uint64_t tail_from_char;   /* getChar */
  uint32_t emulate_tail;     /* 1 = bit-exact with the reference (default), 0 = true rank       */
  /* k-mer suffix table: the interval findInterval holds after consuming the last ftab_k characters of a
   * query (including the value at which it stopped early), indexed by those characters as a base-4
   * number, first character most significant.  {lower, upper} pairs; null / 0 = not built. */
  const uint64_t* ftab;
  uint32_t ftab_k;
  uint32_t sel_shift[RS_NSYM]; /* select directory of c: one entry per 2^sel_shift[c] occurrences */
  /* Locate table, optional (null = not built): where extractPrefix's walk from a row ends, without walking.
   * Every row lies on the walk that starts at the row of some read's terminator (rows 0 .. num_strings-1):
   *   loc[row]   = start_row << 24 | symbols walked from start_row before reaching this row   (2^64-1: none)
   *   loc_fin[s] = terminal row of the walk from start row s << 24 | its length in symbols    (2^64-1: the walk
   *                does not end, reference tail region)
   * so a row's own walk ends at the same terminal after (length - walked) symbols. */
  const uint64_t* loc;
  const uint64_t* loc_fin;
  /* Sampled variant (loc_every > 1): loc holds entries only for the rows reached after a multiple of loc_every
   * symbols of a walk (the start rows included), in row order.  loc_mark = one 16-byte record per 64 rows:
   * {bit i = row 64 w + i is sampled, number of sampled rows before row 64 w}, so the record a walk step fetches to
   * learn whether its row is sampled also gives the slot of the row's entry.  A row's walk is followed until it meets
   * a sampled row or its end: fewer than loc_every LF steps.  n (1/4 + 8/loc_every) bytes instead of 8 n.
   * (loc_rank is unused since the records carry the ranks.) */
  const uint64_t* loc_mark;
  const uint64_t* loc_rank;
  uint32_t loc_every; /* 0: no table, 1: full table, > 1: sampled */
  /* Dense rank blocks, optional (null = not built): the same BWT once more, 128 symbols per 64-byte block as three
   * bit planes, addressed by position alone (block = pos >> 7: no directory, one memory access per rank) and ranked
   * with a handful of logic + popcount instructions instead of a scan over run bytes.  0.5 bytes per symbol: built
   * beside the run blocks when runs are short (a shard of a read-partitioned population: ~3 symbols per run, where
   * the run blocks cost 0.52 bytes per symbol themselves) or when the caller asks for it; rank and access use it,
   * select and the reference's stray-byte rule keep using the run blocks.
   *   +0   u32 cnt[4]   occurrences of A,C,G,T before the block (shards with < 2^32 of each)
   *   +16  u32 p0[4]    bit i of word w: low  bit of (code - 1) of symbol 32 w + i   (0 for '$')
   *   +32  u32 p1[4]                     high bit
   *   +48  u32 p2[4]                     1 = the symbol is '$' */
  const void* dense;
} rs_index_dev;

#define RS_DENSE_SHIFT 7
#define RS_DENSE_SYMS 128u

#endif
