/* Host-side re-encoder: see rs_build.h. */
#include "rs_build.h"

#include <fcntl.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

int rs_read_bwt_file(const char* path, std::vector<uint8_t>& runs, uint64_t& num_strings, uint64_t& num_symbols) {
  /* SGA binary RLE BWT (rlebwt_reader.cpp:27-48; writer submodules/sga/src/SuffixTools/BWTWriterBinary.cpp:28-46):
   * u16 0xCACA, u64 num_strings, u64 num_symbols, u64 num_runs, i32 flag, then num_runs run bytes. */
  FILE* f = fopen(path, "rb");
  if (!f) return RS_ERR_OPEN;
  uint8_t hdr[30];
  if (fread(hdr, 1, sizeof(hdr), f) != sizeof(hdr)) {
    fclose(f);
    return RS_ERR_SHORT;
  }
  uint16_t magic;
  uint64_t nruns;
  memcpy(&magic, hdr, 2);
  memcpy(&num_strings, hdr + 2, 8);
  memcpy(&num_symbols, hdr + 10, 8);
  memcpy(&nruns, hdr + 18, 8);
  if (magic != 0xCACA) {
    fclose(f);
    return RS_ERR_MAGIC;
  }
  /* a corrupt header must not size the allocation: the run bytes are what is left of the file */
  {
    const long at = ftell(f);
    fseek(f, 0, SEEK_END);
    const long end = ftell(f);
    fseek(f, at, SEEK_SET);
    if (at < 0 || end < at || (uint64_t)(end - at) < nruns) {
      fclose(f);
      return RS_ERR_SHORT;
    }
  }
  runs.resize(nruns);
  if (nruns && fread(runs.data(), 1, nruns, f) != nruns) {
    fclose(f);
    return RS_ERR_SHORT;
  }
  fclose(f);
  return RS_OK;
}

namespace {
struct Snap {
  uint64_t tot;
  uint64_t cnt[RS_NSYM];
};
}  // namespace

int rs_build_host_index(const uint8_t* runs, uint64_t R, uint64_t num_strings, uint64_t num_symbols,
                        RsHostIndex& out) {
  if (!runs || R == 0) return RS_ERR_ARG;
  out = RsHostIndex();
  out.num_strings = num_strings;
  out.num_runs = R;

  /* Shape of the tree the reference would build (rlebwt.cpp:51-72): one extra level per factor 1024 of R. */
  uint64_t depth = 1;
  for (uint64_t x = R >> 10; x > 0; x >>= 10) ++depth;
  out.depth = depth;
  const uint64_t shift0 = (depth > 1 ? 4 : 0) + (depth > 2 ? 10 * (depth - 2) : 0);
  const uint64_t top_bucket_runs = (uint64_t)64 << shift0; /* runs per top-level bucket */

  const uint64_t nblocks = (R + RS_PAYLOAD - 1) / RS_PAYLOAD;
  if (nblocks > 0xFFFFFFFFull) return RS_ERR_TOO_LARGE;
  out.blocks.resize(nblocks);

  std::vector<Snap> snaps;
  snaps.reserve(R / top_bucket_runs + 2);

  uint64_t cnt[RS_NSYM] = {0, 0, 0, 0, 0};
  uint64_t tot = 0;
  /* replay of the m_vSum hints (rlebwt.cpp:103-106): value of the last hint and how many there are */
  uint64_t vsum_len = 1, vsum_last = 0, next_sum = 65536;

  uint64_t bi = 0;
  uint32_t fill = RS_PAYLOAD; /* forces a new block at run 0 */
  rs_block* cur = nullptr;
  for (uint64_t i = 0; i < R; ++i) {
    const uint32_t sym = runs[i] >> 5, len = runs[i] & 31u;
    if (sym > 4 || len == 0) return RS_ERR_RUNBYTE;
    if (tot >= next_sum) {
      vsum_last = (i - 1) >> 6;
      ++vsum_len;
      next_sum += 65536;
    }
    if (i % top_bucket_runs == 0) {
      Snap s;
      s.tot = tot;
      memcpy(s.cnt, cnt, sizeof(cnt));
      snaps.push_back(s);
    }
    if (fill == RS_PAYLOAD) {
      cur = &out.blocks[bi++];
      memset(cur, 0, sizeof(*cur));
      cur->start = tot;
      for (int c = 0; c < 4; ++c) {
        if (cnt[c + 1] >> 40) return RS_ERR_TOO_LARGE;
        cur->cnt_lo[c] = (uint32_t)cnt[c + 1];
        cur->cnt_hi |= (uint32_t)(cnt[c + 1] >> 32) << (8 * c);
      }
      fill = 0;
    }
    cur->runs[fill++] = runs[i];
    cnt[sym] += len;
    tot += len;
  }
  if (tot != num_symbols) return RS_ERR_ARG;
  out.n = tot;
  memcpy(out.tot, cnt, sizeof(cnt));
  out.pc[0] = 0;
  for (int c = 1; c < RS_NSYM; ++c) out.pc[c] = out.pc[c - 1] + cnt[c - 1];

  /* Tail behaviour of BPTree::rank / access (BPTree.h:74-94, 135-157; SURVEY.md Q1 + appendix A). */
  out.vlast = vsum_len - 1;
  out.q2_safe = (tot >> 16) <= out.vlast;
  {
    const uint64_t b0 = vsum_last >> shift0;
    const uint64_t next_run = (b0 + 1) * top_bucket_runs;
    uint64_t occ_at_c[RS_NSYM];
    if (next_run < R) {
      const Snap& s = snaps[next_run / top_bucket_runs];
      out.clamp_c = s.tot;
      memcpy(occ_at_c, s.cnt, sizeof(occ_at_c));
    } else {
      out.clamp_c = tot;
      memcpy(occ_at_c, cnt, sizeof(occ_at_c));
    }
    /* BPTree::access (BPTree.h:131-187) in its "same bucket" branch keeps `offset = hint >> shift`
     * at every level while offset is still 0, and only searches below the first non-zero one. */
    {
      uint64_t j = 0, sv = shift0;
      while (j + 1 < depth) {
        if ((vsum_last >> sv) != 0) break;
        ++j;
        const uint64_t rem = depth - j;
        sv = (rem > 1 ? 4 : 0) + (rem > 2 ? 10 * (rem - 2) : 0);
      }
      const uint64_t bucket_runs = (uint64_t)64 << sv;
      const uint64_t next_run_char = ((vsum_last >> sv) + 1) * bucket_runs;
      if (next_run_char == next_run) {
        out.clamp_char = out.clamp_c;
      } else if (next_run_char >= R) {
        out.clamp_char = tot;
      } else {
        uint64_t t = 0;
        for (uint64_t i = 0; i < next_run_char; ++i) t += runs[i] & 31u;
        out.clamp_char = t;
      }
    }
    /* initInterval asks getOcc(b, n-1) (query.cpp:18-21) */
    const bool clamped = (tot >> 16) >= out.vlast && tot - 1 >= out.clamp_c;
    for (int c = 0; c < RS_NSYM; ++c) out.init_occ[c] = clamped ? occ_at_c[c] : cnt[c];
  }

  /* position directory */
  const uint64_t ndir = (tot >> RS_SLOT_SHIFT) + 2;
  out.dir.assign(ndir, 0);
  {
    uint64_t b = 0;
    for (uint64_t j = 0; j < ndir; ++j) {
      const uint64_t p = j << RS_SLOT_SHIFT;
      while (b + 1 < nblocks && out.blocks[b + 1].start <= p) ++b;
      uint64_t e = b;
      int nb = 0;
      uint64_t q = b + 1;
      for (; q < nblocks && out.blocks[q].start < p + RS_SLOT_SIZE; ++q) {
        if (nb < 3) e |= (out.blocks[q].start - p) << (32 + RS_SLOT_SHIFT * nb);
        ++nb;
      }
      if (nb > 3) e |= RS_DIR_MORE;
      out.dir[j] = e;
    }
  }

  /* select directories: about one entry per two blocks for every symbol */
  for (int c = 0; c < RS_NSYM; ++c) {
    uint32_t sh = 0;
    while (sh < RS_SEL_SHIFT_MAX && (cnt[c] >> (sh + 1)) >= nblocks / 2 + 1) ++sh;
    out.sel_shift[c] = sh;
    const uint64_t nent = (cnt[c] >> sh) + 2;
    out.sel[c].assign(nent, (uint32_t)(nblocks - 1));
  }
  {
    uint64_t k[RS_NSYM] = {0, 0, 0, 0, 0};
    for (uint64_t b = 0; b < nblocks; ++b) {
      uint64_t after[RS_NSYM];
      if (b + 1 < nblocks) {
        const rs_block& N = out.blocks[b + 1];
        after[0] = N.start;
        for (int c = 1; c < RS_NSYM; ++c) {
          after[c] = (uint64_t)N.cnt_lo[c - 1] | ((uint64_t)((N.cnt_hi >> (8 * (c - 1))) & 0xFFu) << 32);
          after[0] -= after[c];
        }
      } else {
        memcpy(after, cnt, sizeof(after));
      }
      for (int c = 0; c < RS_NSYM; ++c) {
        while ((k[c] << out.sel_shift[c]) < after[c]) {
          out.sel[c][k[c]] = (uint32_t)b;
          ++k[c];
        }
      }
    }
  }
  return RS_OK;
}

void rs_view_set_tail(const uint64_t init_occ_ref[RS_NSYM], bool on, rs_index_dev* d) {
  d->emulate_tail = on ? 1u : 0u;
  const uint64_t win = d->vlast ? (d->vlast << 16) - 1 : 0; /* (i + 1) >> 16 >= vlast  <=>  i >= win (i != 2^64-1) */
  d->tail_from = on ? (win > d->clamp_c ? win : d->clamp_c) : ~0ull;
  d->tail_from_char = on ? (win > d->clamp_char ? win : d->clamp_char) : ~0ull;
  for (int c = 0; c < RS_NSYM; ++c) d->init_occ[c] = on ? init_occ_ref[c] : d->tot[c];
}

void rs_view_set_tail(const RsHostIndex& h, bool on, rs_index_dev* d) { rs_view_set_tail(h.init_occ, on, d); }

void rs_fill_view(const RsHostIndex& h, rs_index_dev* d) {
  memset(d, 0, sizeof(*d));
  for (int c = 0; c < RS_NSYM; ++c) {
    d->sel_shift[c] = h.sel_shift[c];
    d->pc[c] = h.pc[c];
    d->tot[c] = h.tot[c];
  }
  d->n = h.n;
  d->nblocks = h.blocks.size();
  d->num_strings = h.num_strings;
  d->clamp_c = h.clamp_c;
  d->clamp_char = h.clamp_char;
  d->vlast = h.vlast;
  rs_view_set_tail(h, true, d);
}

namespace {
/* read-only view of a whole file (mmap; an empty file maps to nothing) */
struct MappedFile {
  const char* p = nullptr;
  size_t n = 0;
  int fd = -1;
  bool open(const char* path) {
    fd = ::open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0) return false;
    n = (size_t)st.st_size;
    if (n) {
      void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
      if (m == MAP_FAILED) return false;
      p = (const char*)m;
      madvise(m, n, MADV_SEQUENTIAL);
    }
    return true;
  }
  ~MappedFile() {
    if (p) munmap((void*)p, n);
    if (fd >= 0) ::close(fd);
  }
};

/* next line of [cur, end): returns false at the end; a last line without '\n' counts when non-empty */
inline bool next_line(const char*& cur, const char* end, const char*& b, const char*& e) {
  if (cur >= end) return false;
  b = cur;
  const char* nl = (const char*)memchr(cur, '\n', (size_t)(end - cur));
  e = nl ? nl : end;
  cur = nl ? nl + 1 : end;
  return true;
}
}  // namespace

int rs_parse_etl_samples(const char* path, int sos, int meta, std::vector<uint64_t>& offsets, std::vector<uint8_t>& codes,
                         std::vector<uint8_t>* meta_out) {
  offsets.assign(1, 0);
  codes.clear();
  if (meta_out) meta_out->clear();
  if (sos < 1) return RS_ERR_ARG;
  MappedFile f;
  if (!f.open(path)) return RS_ERR_OPEN;
  const size_t stride = (size_t)sos + (meta ? 2 : 0);
  const char *cur = f.p, *end = f.p + f.n, *rb, *re, *vb, *ve;
  while (next_line(cur, end, rb, re)) {
    if (!next_line(cur, end, vb, ve)) return RS_ERR_ARG; /* a read without its value line */
    if (memchr(rb, 'N', (size_t)(re - rb))) continue;    /* build_bwt.sh:919 drops it before sga index */
    const size_t len = (size_t)(ve - vb);
    if (len % stride != 0) return RS_ERR_ARG;
    for (size_t i = 0; i < len; i += stride) {
      codes.insert(codes.end(), (const uint8_t*)vb + i, (const uint8_t*)vb + i + sos);
      if (meta && meta_out) meta_out->insert(meta_out->end(), (const uint8_t*)vb + i + sos, (const uint8_t*)vb + i + sos + 2);
    }
    offsets.push_back(codes.size() / (size_t)sos);
  }
  return RS_OK;
}

int rs_parse_etl_reads(const char* path, std::string& text, std::vector<uint64_t>& offsets) {
  text.clear();
  offsets.assign(1, 0);
  MappedFile f;
  if (!f.open(path)) return RS_ERR_OPEN;
  const char *cur = f.p, *end = f.p + f.n, *rb, *re, *vb, *ve;
  while (next_line(cur, end, rb, re)) {
    if (!next_line(cur, end, vb, ve)) return RS_ERR_ARG; /* a read without its value line */
    if (memchr(rb, 'N', (size_t)(re - rb))) continue;
    text.append(rb, (size_t)(re - rb));
    offsets.push_back(text.size());
  }
  return RS_OK;
}
