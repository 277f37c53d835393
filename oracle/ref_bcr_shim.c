/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product path.
 *
 * Drives the reference's own BWT builder, Heng Li's BCR as vendored in SGA
 * (submodules/sga/src/Thirdparty/bcr.c, MIT), the way `sga index -a ropebwt
 * --no-reverse` does (submodules/sga/src/SuffixTools/BWTCARopebwt.cpp:29-90):
 * alphabet $ACGTN -> 0..5, bcr_append per read in input order, bcr_build, then
 * iterate the runs and write an SGA binary RLE .bwt
 * (submodules/sga/src/SuffixTools/BWTWriterBinary.cpp:28-93: u16 0xCACA, u64
 * num_strings, u64 num_symbols, u64 num_runs, i32 flag=0, then one byte per
 * run = sym<<5 | len, len<=31, same-symbol runs merged greedily).
 * bcr.c is compiled where it lies by oracle/Makefile; nothing is copied.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "bcr.h"

static uint8_t nt6(char c) {
  switch (c) {
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 3;
    case 'T': case 't': return 4;
    case 0: return 0;
    default: return 5;
  }
}

typedef struct {
  FILE* f;
  int sym;      /* current run symbol (0..4), -1 if none */
  int len;      /* current run length */
  uint64_t nruns;
} rle_writer;

static void rle_flush(rle_writer* w) {
  if (w->sym >= 0 && w->len > 0) {
    uint8_t b = (uint8_t)((w->sym << 5) | w->len);
    fwrite(&b, 1, 1, w->f);
    w->nruns++;
  }
}

static void rle_put(rle_writer* w, int sym, int count) {
  while (count > 0) {
    if (w->sym == sym && w->len < 31) {
      int add = 31 - w->len;
      if (add > count) add = count;
      w->len += add;
      count -= add;
    } else {
      rle_flush(w);
      w->sym = sym;
      w->len = 0;
    }
  }
}

/* reads: concatenated ASCII, read i = reads[off[i] .. off[i+1]).  Returns 0 on success. */
int ref_bcr_build(const char* reads, const uint64_t* off, uint64_t nreads, const char* out_bwt,
                  const char* tmp_path) {
  uint64_t i, nbases = 0;
  bcr_t* b;
  bcritr_t* itr;
  const uint8_t* s;
  int l;
  rle_writer w;
  uint64_t nsym, zero = 0;
  uint16_t magic = 0xCACA;
  int32_t flag = 0;
  long runs_at;

  bcr_verbose = 0;
  b = bcr_init(0, tmp_path);
  for (i = 0; i < nreads; ++i) {
    uint64_t len = off[i + 1] - off[i], j;
    uint8_t* enc = (uint8_t*)malloc(len ? len : 1);
    for (j = 0; j < len; ++j) {
      enc[j] = nt6(reads[off[i] + j]);
      if (enc[j] == 0 || enc[j] == 5) { free(enc); bcr_destroy(b); return 2; } /* pipeline drops N reads */
    }
    bcr_append(b, (int)len, enc);
    free(enc);
    nbases += len;
  }
  bcr_build(b);

  w.f = fopen(out_bwt, "wb");
  if (!w.f) { bcr_destroy(b); return 1; }
  w.sym = -1; w.len = 0; w.nruns = 0;
  nsym = nbases + nreads;
  fwrite(&magic, 2, 1, w.f);
  fwrite(&nreads, 8, 1, w.f);
  fwrite(&nsym, 8, 1, w.f);
  runs_at = ftell(w.f);
  fwrite(&zero, 8, 1, w.f);
  fwrite(&flag, 4, 1, w.f);

  itr = bcr_itr_init(b);
  while ((s = bcr_itr_next(itr, &l)) != 0) {
    int k;
    for (k = 0; k < l; ++k) rle_put(&w, s[k] & 7, s[k] >> 3);
  }
  free(itr);
  rle_flush(&w);
  fseek(w.f, runs_at, SEEK_SET);
  fwrite(&w.nruns, 8, 1, w.f);
  fclose(w.f);

  bcr_destroy(b);
  return 0;
}
