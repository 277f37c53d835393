/* TEST INFRASTRUCTURE ONLY -- see rs_oracle.h.  Plain-C restatement of the reference algorithm.
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 * Deliberately keeps the reference's structure (multi-level BP-tree sampled on run index, binary
 * search on cumulative sums, m_vSum hints) INCLUDING its tail behaviour: that
 * behaviour must emerge here from the algorithm, not from a special case.
 */
#define _POSIX_C_SOURCE 200809L
#include "rs_oracle.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define SAMPLE_SMALL 64u      /* rlebwt.h:11 DEFAULT_SAMPLE_RATE_SMALL */
#define SAMPLE_LARGE 1024u    /* rlebwt.h:12 DEFAULT_SAMPLE_RATE_LARGE */
#define MAX_RUN 31u           /* rlebwt.h:13 DEFAULT_MAX_COUNT */
#define SYMBOL_RATE 65536u    /* BPTree.h:9  DEFAULT_SYMBOL_RATE */
#define SYMBOL_SHIFT 16       /* BPTree.h:10 */
#define LARGE_SHIFT 10        /* BPTree.h:11 */
#define SMALL_SHIFT 4         /* BPTree.h:12 */
#define NSYM 5

static const char RANK_ALPHABET[NSYM] = {'$', 'A', 'C', 'G', 'T'}; /* alphabet.h:9 */

/* alphabet.h:11-32: every byte that is not A,C,G,T ranks as '$' (0). */
static inline unsigned get_rank(char b) {
  switch (b) {
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 3;
    case 'T': return 4;
    default: return 0;
  }
}

/* One BPNodes<IntType> level (BPNodes.h:34-241). Values are kept widened to u64; `width` remembers
 * sizeof(IntType) for the .bpi2 format. No level can overflow its IntType (rlebwt.cpp:55-70 picks it so). */
typedef struct {
  uint64_t width;       /* 2, 4 or 8 */
  uint64_t length;      /* m_length */
  uint64_t block_size;  /* m_block_size: buckets per parent bucket */
  uint64_t bucket_size; /* m_bucket_size: runs per bucket */
  uint64_t cap;
  uint64_t* counts;     /* length x 5 (m_array) */
  uint64_t* sums;       /* length     (m_sum)   */
  uint64_t last[NSYM];  /* m_last */
} level_t;

struct rso_index {
  uint64_t num_strings, num_symbols, num_runs;
  uint8_t* runs;
  uint64_t depth;
  level_t* lv;
  uint32_t* vsum;
  uint64_t vsum_len, vsum_cap;
  uint64_t pred[NSYM]; /* m_predCount */
  uint64_t q2;         /* out-of-bounds m_vSum reads that the reference would have made */
};

/* ---------------- level helpers ---------------- */
static void level_init(level_t* L, uint64_t width, uint64_t reserve, uint64_t bucket, uint64_t block) {
  memset(L, 0, sizeof(*L));
  L->width = width;
  L->bucket_size = bucket;
  L->block_size = block;
  L->cap = reserve ? reserve : 1;
  L->counts = (uint64_t*)malloc(L->cap * NSYM * sizeof(uint64_t));
  L->sums = (uint64_t*)malloc(L->cap * sizeof(uint64_t));
}

/* BPNodes.h:92-96 appendLast */
static void level_append_last(level_t* L) {
  if (L->length == L->cap) {
    L->cap *= 2;
    L->counts = (uint64_t*)realloc(L->counts, L->cap * NSYM * sizeof(uint64_t));
    L->sums = (uint64_t*)realloc(L->sums, L->cap * sizeof(uint64_t));
  }
  uint64_t s = 0;
  for (int c = 0; c < NSYM; ++c) {
    L->counts[L->length * NSYM + c] = L->last[c];
    s += L->last[c];
  }
  L->sums[L->length] = s; /* alphabet.h:73-80 getSum */
  L->length++;
}

/* BPNodes.h:133-162 get_access: last p in [lower,upper] with sum[p] < cnt, else lower. */
static uint64_t level_get_access(const level_t* L, uint64_t cnt, uint64_t lower, uint64_t upper) {
  uint64_t begin = lower, end = upper;
  if (end > L->length - 1) end = L->length - 1;
  while (end > begin) {
    uint64_t p = (begin + end + 1) >> 1;
    if (L->sums[p] >= cnt) {
      if (end == p) break;
      end = p;
    } else {
      if (begin == p) break;
      begin = p;
    }
  }
  return begin;
}

/* BPNodes.h:118-124 access(cnt, lower, upper) */
static uint64_t level_access_range(const level_t* L, uint64_t cnt, uint64_t lower, uint64_t upper) {
  if (lower == upper) return lower;
  return level_get_access(L, cnt, lower, upper);
}

/* BPNodes.h:126-131 access(cnt, offset) */
static uint64_t level_access_block(const level_t* L, uint64_t cnt, uint64_t offset) {
  uint64_t begin = L->block_size * offset;
  uint64_t end = L->block_size + begin - 1;
  return level_get_access(L, cnt, begin, end);
}

/* BPNodes.h:164-196 select */
static uint64_t level_select(const level_t* L, unsigned sym, uint64_t cnt, uint64_t offset) {
  uint64_t begin = L->block_size * offset;
  uint64_t end = L->block_size + begin - 1;
  if (cnt == 0) return begin; /* `cnt <= 0` on an unsigned */
  if (end > L->length - 1) end = L->length - 1;
  while (end > begin) {
    uint64_t p = (begin + end + 1) >> 1;
    if (L->counts[p * NSYM + sym] >= cnt) {
      if (end == p) break;
      end = p;
    } else {
      if (begin == p) break;
      begin = p;
    }
  }
  return begin;
}

/* ---------------- construction ---------------- */
/* rlebwt.cpp:34-148 initialiseFMIndex */
static void build_fm_index(rso_index* ix) {
  const uint64_t lsv = 10, ssv = 6, sv = SYMBOL_SHIFT;
  const uint64_t R = ix->num_runs;
  /* upper levels, built bottom-up and inserted at the front (rlebwt.cpp:51-72) */
  uint64_t nupper = 0;
  for (uint64_t x = R >> lsv; x > 0; x >>= lsv) ++nupper;
  ix->depth = nupper + 1;
  ix->lv = (level_t*)calloc(ix->depth, sizeof(level_t));
  {
    uint64_t nb = R >> lsv, chars = 1, pos = nupper;
    while (nb > 0) {
      chars *= SAMPLE_LARGE;
      uint64_t maxc = chars * SAMPLE_LARGE * MAX_RUN;
      uint64_t width;
      maxc >>= 16;
      if (maxc < 1) width = 2;
      else {
        maxc >>= 16;
        width = (maxc < 1) ? 4 : 8;
      }
      --pos;
      level_init(&ix->lv[pos], width, nb + 1, chars, SAMPLE_LARGE);
      nb >>= lsv;
    }
  }
  /* bottom level (rlebwt.cpp:74-78) */
  level_init(&ix->lv[ix->depth - 1], 2, (R >> ssv) + 1, SAMPLE_SMALL, (uint64_t)1 << (lsv - ssv));

  ix->vsum_cap = (ix->num_symbols >> sv) + 2;
  ix->vsum = (uint32_t*)malloc(ix->vsum_cap * sizeof(uint32_t));
  ix->vsum_len = 0;
  ix->vsum[ix->vsum_len++] = 0;
  uint64_t next_sum = SYMBOL_RATE, running_total = 0;

  const uint64_t depth = ix->depth, bottom = depth - 1;
  uint64_t* next_buckets = (uint64_t*)calloc(depth, sizeof(uint64_t));
  level_t* lv = ix->lv;

  for (uint64_t i = 0; i < R; ++i) {
    const unsigned sym = ix->runs[i] >> 5;
    const unsigned run_len = ix->runs[i] & 31u;
    if (running_total >= next_sum) { /* rlebwt.cpp:103-106 */
      if (ix->vsum_len == ix->vsum_cap) {
        ix->vsum_cap *= 2;
        ix->vsum = (uint32_t*)realloc(ix->vsum, ix->vsum_cap * sizeof(uint32_t));
      }
      ix->vsum[ix->vsum_len++] = (uint32_t)(lv[bottom].length - 1);
      next_sum += SYMBOL_RATE;
    }
    if (i == next_buckets[bottom]) { /* rlebwt.cpp:108-124 */
      for (uint64_t j = 0; j < depth; ++j) {
        if (i == next_buckets[j]) {
          for (uint64_t k = bottom; k > j; --k) {
            const uint64_t l = k - 1;
            for (int c = 0; c < NSYM; ++c) lv[l].last[c] += lv[k].last[c];
            memset(lv[k].last, 0, sizeof(lv[k].last));
            level_append_last(&lv[k]);
            next_buckets[k] = lv[k].bucket_size * lv[k].length;
          }
          level_append_last(&lv[j]);
          next_buckets[j] = lv[j].bucket_size * lv[j].length;
          break;
        }
      }
    }
    lv[bottom].last[sym] += run_len; /* rlebwt.cpp:126 */
    running_total += run_len;
  }
  free(next_buckets);

  /* rlebwt.cpp:129-147 predCount from the per-level running totals */
  uint64_t tot[NSYM] = {0, 0, 0, 0, 0};
  for (uint64_t d = 0; d < depth; ++d)
    for (int c = 0; c < NSYM; ++c) tot[c] += lv[d].last[c];
  ix->pred[0] = 0;
  for (int c = 1; c < NSYM; ++c) ix->pred[c] = ix->pred[c - 1] + tot[c - 1];
}

static int check_runs(const uint8_t* runs, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i)
    if ((runs[i] >> 5) > 4 || (runs[i] & 31u) == 0) return 4;
  return 0;
}

int rso_from_runs(const uint8_t* runs, uint64_t nruns, uint64_t num_strings, uint64_t num_symbols,
                  rso_index** out) {
  int rc = check_runs(runs, nruns);
  if (rc) return rc;
  rso_index* ix = (rso_index*)calloc(1, sizeof(rso_index));
  ix->num_strings = num_strings;
  ix->num_symbols = num_symbols;
  ix->num_runs = nruns;
  ix->runs = (uint8_t*)malloc(nruns ? nruns : 1);
  memcpy(ix->runs, runs, nruns);
  build_fm_index(ix);
  *out = ix;
  return 0;
}

static int read_u64(FILE* f, uint64_t* v) { return fread(v, 8, 1, f) == 1 ? 0 : 3; }

/* rlebwt.cpp:163-200 deserialiseFMIndex + BPNodes.h:47-66 + alphabet.h:47-49 */
static int load_bpi2(rso_index* ix, const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) return 1;
  uint64_t depth = 0;
  if (read_u64(f, &depth)) { fclose(f); return 3; }
  ix->depth = depth;
  ix->lv = (level_t*)calloc(depth, sizeof(level_t));
  for (uint64_t d = 0; d < depth; ++d) {
    level_t* L = &ix->lv[d];
    uint64_t width, length, block, bucket;
    if (read_u64(f, &width) || read_u64(f, &length) || read_u64(f, &block) || read_u64(f, &bucket)) {
      fclose(f);
      return 3;
    }
    level_init(L, width, length, bucket, block);
    L->length = length;
    for (uint64_t i = 0; i < length * NSYM; ++i) {
      uint64_t v = 0;
      if (fread(&v, width, 1, f) != 1) { fclose(f); return 3; }
      L->counts[i] = v;
    }
    for (uint64_t i = 0; i < length; ++i) {
      uint64_t v = 0;
      if (fread(&v, width, 1, f) != 1) { fclose(f); return 3; }
      L->sums[i] = v;
    }
  }
  uint64_t sz = 0;
  if (read_u64(f, &sz)) { fclose(f); return 3; }
  ix->vsum_len = ix->vsum_cap = sz;
  ix->vsum = (uint32_t*)malloc((sz ? sz : 1) * sizeof(uint32_t));
  if (sz && fread(ix->vsum, 4, sz, f) != sz) { fclose(f); return 3; }
  if (fread(ix->pred, 8, NSYM, f) != NSYM) { fclose(f); return 3; }
  fclose(f);
  return 0;
}

int rso_open(const char* bwt_path, int use_bpi2, rso_index** out) {
  FILE* f = fopen(bwt_path, "rb");
  if (!f) return 1;
  uint16_t magic = 0;
  uint64_t nstr = 0, nsym = 0, nruns = 0;
  int32_t flag = 0;
  /* rlebwt_reader.cpp:27-48: u16 magic 0xCACA, u64 x3, i32 flag */
  if (fread(&magic, 2, 1, f) != 1) { fclose(f); return 3; }
  if (magic != 0xCACA) { fclose(f); return 2; }
  if (read_u64(f, &nstr) || read_u64(f, &nsym) || read_u64(f, &nruns) || fread(&flag, 4, 1, f) != 1) {
    fclose(f);
    return 3;
  }
  uint8_t* runs = (uint8_t*)malloc(nruns ? nruns : 1);
  if (nruns && fread(runs, 1, nruns, f) != nruns) { free(runs); fclose(f); return 3; }
  fclose(f);
  int rc = check_runs(runs, nruns);
  if (rc) { free(runs); return rc; }
  rso_index* ix = (rso_index*)calloc(1, sizeof(rso_index));
  ix->num_strings = nstr;
  ix->num_symbols = nsym;
  ix->num_runs = nruns;
  ix->runs = runs;
  int loaded = 0;
  if (use_bpi2) {
    size_t L = strlen(bwt_path);
    char* p = (char*)malloc(L + 6);
    memcpy(p, bwt_path, L);
    memcpy(p + L, ".bpi2", 6);
    FILE* t = fopen(p, "rb");
    if (t) {
      fclose(t);
      rc = load_bpi2(ix, p);
      if (rc) { free(p); rso_close(ix); return rc; }
      loaded = 1;
    }
    free(p);
  }
  if (!loaded) build_fm_index(ix);
  *out = ix;
  return 0;
}

void rso_close(rso_index* ix) {
  if (!ix) return;
  for (uint64_t d = 0; d < ix->depth; ++d) {
    free(ix->lv[d].counts);
    free(ix->lv[d].sums);
  }
  free(ix->lv);
  free(ix->vsum);
  free(ix->runs);
  free(ix);
}

int rso_write_bpi2(const rso_index* ix, const char* path) {
  FILE* f = fopen(path, "wb");
  if (!f) return 1;
  fwrite(&ix->depth, 8, 1, f); /* rlebwt.cpp:154-155 */
  for (uint64_t d = 0; d < ix->depth; ++d) { /* BPNodes.h:198-214 */
    const level_t* L = &ix->lv[d];
    fwrite(&L->width, 8, 1, f);
    fwrite(&L->length, 8, 1, f);
    fwrite(&L->block_size, 8, 1, f);
    fwrite(&L->bucket_size, 8, 1, f);
    for (uint64_t i = 0; i < L->length * NSYM; ++i) fwrite(&L->counts[i], L->width, 1, f);
    for (uint64_t i = 0; i < L->length; ++i) fwrite(&L->sums[i], L->width, 1, f);
  }
  fwrite(&ix->vsum_len, 8, 1, f); /* BPTree.h:194-198 */
  fwrite(ix->vsum, 4, ix->vsum_len, f);
  fwrite(ix->pred, 8, NSYM, f); /* rlebwt.cpp:159 */
  fclose(f);
  return 0;
}

uint64_t rso_bwlen(const rso_index* ix) { return ix->num_symbols; }
uint64_t rso_num_strings(const rso_index* ix) { return ix->num_strings; }
uint64_t rso_num_runs(const rso_index* ix) { return ix->num_runs; }
uint64_t rso_depth(const rso_index* ix) { return ix->depth; }
uint64_t rso_vsum_size(const rso_index* ix) { return ix->vsum_len; }
uint64_t rso_q2_violations(const rso_index* ix) { return ix->q2; }

/* ---------------- BPTree: rank / access / select ---------------- */
typedef struct {
  uint64_t index, sum, count;
} marker_t;

/* m_vSum[lb] with the Q2 guard (the reference reads out of bounds instead). */
static inline uint64_t vsum_at(const rso_index* ix, uint64_t lb) {
  if (lb >= ix->vsum_len) {
    ((rso_index*)ix)->q2++;
    return ix->vsum[ix->vsum_len - 1];
  }
  return ix->vsum[lb];
}

static inline uint64_t level_shift(uint64_t td, uint64_t i) { /* BPTree.h:97, 147 */
  uint64_t rem = td - i;
  return SMALL_SHIFT * (rem > 1 ? 1 : 0) + LARGE_SHIFT * (rem > 2 ? rem - 2 : 0);
}

/* BPTree.h:69-129 */
static marker_t tree_rank(const rso_index* ix, unsigned sym, uint64_t index) {
  uint64_t offset = 0, count = 0, sum = 0;
  uint64_t lb = index >> SYMBOL_SHIFT;
  const uint64_t idx_at = vsum_at(ix, lb);
  uint64_t next_at = idx_at;
  ++lb;
  if (lb < ix->vsum_len) next_at = ix->vsum[lb];
  const uint64_t td = ix->depth;
  const level_t* lv = ix->lv;

  if (idx_at == next_at) {
    offset = idx_at >> (SMALL_SHIFT * (td > 1 ? 1 : 0) + LARGE_SHIFT * (td > 2 ? td - 2 : 0));
    sum += lv[0].sums[offset];
    count += lv[0].counts[offset * NSYM + sym];
    for (uint64_t i = 1; i < td; ++i) {
      offset = level_access_block(&lv[i], index - sum, offset);
      sum += lv[i].sums[offset];
      count += lv[i].counts[offset * NSYM + sym];
    }
  } else {
    for (uint64_t i = 0; i < td; ++i) {
      const uint64_t sv = level_shift(td, i);
      uint64_t lower = idx_at >> sv, upper = next_at >> sv;
      const uint64_t bs = lv[i].block_size;
      const uint64_t lower_bound = offset * bs, upper_bound = lower_bound + bs - 1;
      if (lower_bound > lower) lower = lower_bound;
      if (upper_bound < upper) upper = upper_bound;
      offset = level_access_range(&lv[i], index - sum, lower, upper);
      sum += lv[i].sums[offset];
      count += lv[i].counts[offset * NSYM + sym];
    }
  }
  if (offset == 0) offset = idx_at;
  marker_t mk = {offset, sum, count};
  return mk;
}

/* BPTree.h:131-187 */
static marker_t tree_access(const rso_index* ix, uint64_t index) {
  uint64_t offset = 0, sum = 0;
  uint64_t lb = index >> SYMBOL_SHIFT;
  const uint64_t idx_at = vsum_at(ix, lb);
  uint64_t next_at = idx_at;
  ++lb;
  if (lb < ix->vsum_len) next_at = ix->vsum[lb];
  const uint64_t td = ix->depth;
  for (uint64_t i = 0; i < td; ++i) {
    const level_t* L = &ix->lv[i];
    const uint64_t sv = level_shift(td, i);
    uint64_t lower = idx_at >> sv;
    if (idx_at == next_at) {
      if (offset == 0) offset = lower;
      else offset = level_access_block(L, index - sum, offset);
    } else {
      uint64_t upper = next_at >> sv;
      const uint64_t bs = L->block_size;
      const uint64_t lower_bound = offset * bs, upper_bound = lower_bound + bs - 1;
      if (lower_bound > lower) lower = lower_bound;
      if (upper_bound < upper) upper = upper_bound;
      offset = level_access_range(L, index - sum, lower, upper);
    }
    sum += L->sums[offset];
  }
  if (offset == 0) offset = idx_at;
  marker_t mk = {offset, sum, sum};
  return mk;
}

/* BPTree.h:50-67 */
static marker_t tree_select(const rso_index* ix, unsigned sym, uint64_t bc) {
  uint64_t offset = 0, count = 0, sum = 0;
  for (uint64_t i = 0; i < ix->depth; ++i) {
    const level_t* L = &ix->lv[i];
    offset = level_select(L, sym, bc - count, offset);
    sum += L->sums[offset];
    count += L->counts[offset * NSYM + sym];
  }
  marker_t mk = {offset, sum, count};
  return mk;
}

/* ---------------- RLEBWT getters ---------------- */
uint64_t rso_get_pc(const rso_index* ix, char b) { return ix->pred[get_rank(b)]; }

uint64_t rso_get_occ(const rso_index* ix, char b, uint64_t index) {
  const unsigned sym = get_rank(b);
  const uint64_t idx = index + 1;
  const marker_t mk = tree_rank(ix, sym, idx);
  const uint64_t begin = mk.index * SAMPLE_SMALL;
  uint64_t occ = mk.count, offset = idx - mk.sum;
  const uint64_t length = ix->num_runs;
  uint64_t end = begin + SAMPLE_SMALL - 1;
  if (end >= length) end = length - 1;
  for (uint64_t i = begin; i <= end; ++i) {
    const uint64_t count = ix->runs[i] & 31u;
    /* rlebwt.cpp:286,293 compare CHARACTERS (b == unit.getChar()): a byte outside "$ACGT" takes the
     * '$' counters in the tree (alphabet.h getRank) but matches no run here. */
    const char rc = RANK_ALPHABET[ix->runs[i] >> 5];
    if (offset <= count) {
      if (b == rc) occ += offset;
      break;
    } else {
      offset -= count;
      if (b == rc) occ += count;
    }
  }
  return occ;
}

uint64_t rso_get_occ_at(const rso_index* ix, char b, uint64_t bc) {
  const unsigned sym = get_rank(b);
  const marker_t mk = tree_select(ix, sym, bc);
  const uint64_t begin = mk.index * SAMPLE_SMALL;
  uint64_t end = begin + SAMPLE_SMALL - 1;
  uint64_t offset = bc - mk.count;
  const uint64_t length = ix->num_runs;
  if (end >= length) end = length - 1;
  uint64_t index = mk.sum;
  for (uint64_t i = begin; i <= end; ++i) {
    const uint64_t count = ix->runs[i] & 31u;
    if (b != RANK_ALPHABET[ix->runs[i] >> 5]) { /* rlebwt.cpp:251 compares characters */
      index += count;
      continue;
    }
    if (offset <= count) {
      index += offset - 1;
      break;
    } else {
      offset -= count;
      index += count;
    }
  }
  return index;
}

char rso_get_char(const rso_index* ix, uint64_t index) {
  const uint64_t idx = index + 1;
  const marker_t mk = tree_access(ix, idx);
  const uint64_t begin = mk.index * SAMPLE_SMALL;
  uint64_t end = begin + SAMPLE_SMALL - 1;
  uint64_t offset = idx - mk.sum;
  const uint64_t length = ix->num_runs;
  if (end >= length) end = length - 1;
  for (uint64_t i = begin; i <= end; ++i) {
    const uint64_t count = ix->runs[i] & 31u;
    if (offset <= count) return RANK_ALPHABET[ix->runs[i] >> 5];
    offset -= count;
  }
  return RANK_ALPHABET[ix->runs[end] >> 5];
}

char rso_get_f(const rso_index* ix, uint64_t index) {
  unsigned ci = 0;
  while (ci < NSYM && ix->pred[ci] <= index) ci++;
  return RANK_ALPHABET[ci - 1];
}

/* ---------------- query.cpp ---------------- */
void rso_find_interval(const rso_index* ix, const char* w, size_t len, uint64_t* lower_out,
                       uint64_t* upper_out) {
  long j = (long)len - 1;
  char curr = w[j];
  /* initInterval, query.cpp:18-21 */
  uint64_t lower = rso_get_pc(ix, curr);
  uint64_t upper = lower + rso_get_occ(ix, curr, ix->num_symbols - 1) - 1;
  --j;
  for (; j >= 0; --j) {
    curr = w[j];
    /* updateInterval, query.cpp:11-15 */
    const uint64_t pb = rso_get_pc(ix, curr);
    lower = pb + rso_get_occ(ix, curr, lower - 1);
    upper = pb + rso_get_occ(ix, curr, upper) - 1;
    if (lower > upper) break; /* query.cpp:35-37 */
  }
  *lower_out = lower;
  *upper_out = upper;
}

size_t rso_extract_prefix(const rso_index* ix, uint64_t row, char* out, size_t cap, uint64_t* terminal,
                          size_t max_steps) {
  uint64_t idx = row;
  size_t n = 0;
  while (1) {
    char b = rso_get_char(ix, idx);
    idx = rso_get_pc(ix, b) + rso_get_occ(ix, b, idx - 1);
    if (b == '$') break;
    if (out && n + 1 < cap) out[n] = b;
    ++n;
    if (n > max_steps) { /* the reference would never return (walk cycling inside the Q1 tail) */
      if (terminal) *terminal = UINT64_MAX;
      if (out && cap) out[0] = 0;
      return (size_t)-1;
    }
  }
  if (out && cap) {
    size_t m = n < cap - 1 ? n : cap - 1;
    for (size_t i = 0; i < m / 2; ++i) { /* std::reverse, query.cpp:60 */
      char t = out[i];
      out[i] = out[m - 1 - i];
      out[m - 1 - i] = t;
    }
    out[m] = 0;
  }
  if (terminal) *terminal = idx;
  return n;
}

size_t rso_extract_postfix(const rso_index* ix, uint64_t row, char* out, size_t cap, uint64_t* terminal,
                           size_t max_steps) {
  uint64_t idx = row;
  size_t n = 0;
  while (1) {
    char f = rso_get_f(ix, idx);
    uint64_t fc = idx - rso_get_pc(ix, f) + 1;
    const uint64_t at = idx; /* row in F: when f == '$' this row IS the read's input-order index */
    idx = rso_get_occ_at(ix, f, fc);
    if (f == '$') {
      idx = at;
      break;
    }
    if (out && n + 1 < cap) out[n] = f;
    ++n;
    if (n > max_steps) {
      if (terminal) *terminal = UINT64_MAX;
      if (out && cap) out[0] = 0;
      return (size_t)-1;
    }
  }
  if (out && cap) out[n < cap - 1 ? n : cap - 1] = 0;
  if (terminal) *terminal = idx;
  return n;
}

/* ---------------- batch drivers ---------------- */
static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

double rso_get_occ_batch(const rso_index* ix, const char* b, const uint64_t* idx, size_t n, uint64_t* out) {
  double t0 = now_s();
  for (size_t i = 0; i < n; ++i) out[i] = rso_get_occ(ix, b[i], idx[i]);
  return now_s() - t0;
}

double rso_get_char_batch(const rso_index* ix, const uint64_t* idx, size_t n, char* out) {
  double t0 = now_s();
  for (size_t i = 0; i < n; ++i) out[i] = rso_get_char(ix, idx[i]);
  return now_s() - t0;
}

double rso_get_occ_at_batch(const rso_index* ix, const char* b, const uint64_t* cnt, size_t n, uint64_t* out) {
  double t0 = now_s();
  for (size_t i = 0; i < n; ++i) out[i] = rso_get_occ_at(ix, b[i], cnt[i]);
  return now_s() - t0;
}

typedef struct {
  const rso_index* ix;
  size_t t, n, stride;
  /* find_interval */
  const char* kmers;
  int k;
  uint64_t *lower, *upper;
  /* locate */
  const uint64_t* rows;
  uint64_t *lexrank, *read_index;
  uint32_t *pre_steps, *post_steps;
  size_t max_steps;
} job_t;

static void* fi_worker(void* p) {
  job_t* j = (job_t*)p;
  for (size_t q = j->t; q < j->n; q += j->stride)
    rso_find_interval(j->ix, j->kmers + q * (size_t)j->k, (size_t)j->k, &j->lower[q], &j->upper[q]);
  return NULL;
}

static void* loc_worker(void* p) {
  job_t* j = (job_t*)p;
  for (size_t i = j->t; i < j->n; i += j->stride) {
    uint64_t term = 0;
    if (j->lexrank) {
      size_t s = rso_extract_prefix(j->ix, j->rows[i], NULL, 0, &term, j->max_steps);
      j->lexrank[i] = term;
      if (j->pre_steps) j->pre_steps[i] = (uint32_t)s;
    }
    if (j->read_index) {
      size_t s = rso_extract_postfix(j->ix, j->rows[i], NULL, 0, &term, j->max_steps);
      j->read_index[i] = term;
      if (j->post_steps) j->post_steps[i] = (uint32_t)s;
    }
  }
  return NULL;
}

static double run_jobs(job_t proto, int nthreads, void* (*fn)(void*)) {
  if (nthreads < 1) nthreads = 1;
  double t0 = now_s();
  if (nthreads == 1) {
    proto.t = 0;
    proto.stride = 1;
    fn(&proto);
  } else {
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    job_t* jobs = (job_t*)malloc(sizeof(job_t) * (size_t)nthreads);
    for (int t = 0; t < nthreads; ++t) {
      jobs[t] = proto;
      jobs[t].t = (size_t)t;
      jobs[t].stride = (size_t)nthreads;
      pthread_create(&th[t], NULL, fn, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
  }
  return now_s() - t0;
}

double rso_find_interval_batch(const rso_index* ix, const char* kmers, int k, size_t Q, uint64_t* lower,
                               uint64_t* upper, int nthreads) {
  job_t j;
  memset(&j, 0, sizeof(j));
  j.ix = ix; j.n = Q; j.kmers = kmers; j.k = k; j.lower = lower; j.upper = upper;
  return run_jobs(j, nthreads, fi_worker);
}

double rso_locate_batch(const rso_index* ix, const uint64_t* rows, size_t n, uint64_t* lexrank,
                        uint32_t* pre_steps, uint64_t* read_index, uint32_t* post_steps, int nthreads,
                        uint32_t max_steps) {
  job_t j;
  memset(&j, 0, sizeof(j));
  j.ix = ix; j.n = n; j.rows = rows; j.lexrank = lexrank; j.read_index = read_index;
  j.pre_steps = pre_steps; j.post_steps = post_steps; j.max_steps = max_steps;
  return run_jobs(j, nthreads, loc_worker);
}

/* ---------------- sample store value format ---------------- */
size_t rso_decode_sample_value(const uint8_t* value, size_t value_len, int size_of_sample,
                               int has_other_meta_data, uint8_t* codes_out, int32_t* counts,
                               int32_t* lanes, size_t cap) {
  /* service.cpp:829-843 */
  size_t pos = 0, n = 0;
  while (pos < value_len) {
    if (n < cap && codes_out) {
      for (int i = 0; i < size_of_sample; ++i)
        codes_out[n * (size_t)size_of_sample + (size_t)i] = pos + (size_t)i < value_len ? value[pos + (size_t)i] : 0;
    }
    pos += (size_t)size_of_sample;
    int32_t c = 0, l = 0;
    if (has_other_meta_data) {
      c = pos < value_len ? (int32_t)(int8_t)value[pos] - 33 : -33;
      ++pos;
      l = pos < value_len ? (int32_t)(int8_t)value[pos] - 33 : -33;
      ++pos;
    }
    if (n < cap) {
      if (counts) counts[n] = c;
      if (lanes) lanes[n] = l;
    }
    ++n;
  }
  return n;
}

void rso_sample_code(uint64_t sample_idx, int size_of_sample, uint8_t* out) {
  /* create_samples_hash.cpp:41-50: curr_value is 1-based; digits base 223, offset 33, MSB first */
  uint64_t rem = sample_idx + 1;
  for (int i = 1; i <= size_of_sample; ++i) {
    out[size_of_sample - i] = (uint8_t)(rem % 223 + 33);
    rem /= 223;
  }
}
