"""Synthetic population-BWT inputs of the shapes BASELINE.json names.

This is an input generator, not part of the query path: it makes genomes, reads, the SGA-format
run-length BWT (``.bwt``) of a read set, and query batches.  Everything is plain torch tensor code so
the same functions run on the CPU for the small parity cases and on the GPU for the full-size
configurations (a 1.1 G-symbol BWT is a handful of 64-bit radix sorts on a B200).

Conventions follow the reference's ETL:
* reads are de-duplicated and fed to the BWT builder in RLO order -- sorted by *reversed* string
  (``src/util/rlosort_seq_and_convert_sample_names.cpp:16-20``; ``demo/build_bwt.sh`` steps 6-8);
* the BWT is the one ``sga index -a ropebwt --no-reverse`` produces: suffixes sorted with
  ``$ < A < C < G < T`` and ``$_i < $_j  <=>  i < j`` (probe 1), written as
  ``BWTWriterBinary.cpp:28-93`` does: 30-byte header then one byte ``sym<<5 | len`` per run, ``len <= 31``.

Symbols are coded 0..4 = ``$ACGT`` throughout (``include/bwt/alphabet.h:9``).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np
import torch

CHARS_PER_WORD = 21  # 21 symbols x 3 bits = 63 bits: keys stay non-negative in int64
ASCII = np.frombuffer(b"$ACGT", dtype=np.uint8)
BWT_MAGIC = 0xCACA
BWT_HEADER_BYTES = 30


# ----------------------------------------------------------------------------------------------
# genomes / reads / queries
# ----------------------------------------------------------------------------------------------
def _gen(seed: int, device) -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def random_genome(length: int, seed: int, device="cpu") -> torch.Tensor:
    """Uniform random ACGT, codes 1..4, uint8."""
    return torch.randint(1, 5, (length,), generator=_gen(seed, device), device=device, dtype=torch.uint8)


def mutate_snps(genome: torch.Tensor, rate: float, seed: int, sites: torch.Tensor | None = None) -> torch.Tensor:
    """Copy of `genome` with substitutions at ~`rate` of positions (or at `sites` drawn with prob. rate)."""
    dev = genome.device
    g = _gen(seed, dev)
    out = genome.clone()
    if sites is None:
        hit = torch.rand(genome.numel(), generator=g, device=dev) < rate
    else:
        pick = torch.rand(sites.numel(), generator=g, device=dev) < rate
        hit = torch.zeros(genome.numel(), dtype=torch.bool, device=dev)
        hit[sites[pick]] = True
    shift = torch.randint(1, 4, (genome.numel(),), generator=g, device=dev, dtype=torch.uint8)
    alt = (genome - 1 + shift) % 4 + 1
    out[hit] = alt[hit]
    return out


def sample_reads(haplotypes: list[torch.Tensor], n_reads: int, read_len: int, seed: int,
                 min_len: int | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """n_reads forward-strand, error-free reads, uniform start, random haplotype.
    Returns (reads [n, read_len] uint8 zero padded, lengths [n] int32)."""
    dev = haplotypes[0].device
    g = _gen(seed, dev)
    G = haplotypes[0].numel()
    starts = torch.randint(0, G - read_len + 1, (n_reads,), generator=g, device=dev)
    hap = torch.randint(0, len(haplotypes), (n_reads,), generator=g, device=dev)
    H = torch.stack(haplotypes)  # [h, G]
    cols = torch.arange(read_len, device=dev)
    reads = torch.empty((n_reads, read_len), dtype=torch.uint8, device=dev)
    chunk = 1 << 22
    for s in range(0, n_reads, chunk):
        e = min(n_reads, s + chunk)
        idx = hap[s:e, None] * G + starts[s:e, None] + cols[None, :]
        reads[s:e] = H.reshape(-1)[idx]
    if min_len is None or min_len >= read_len:
        lengths = torch.full((n_reads,), read_len, dtype=torch.int32, device=dev)
    else:
        lengths = torch.randint(min_len, read_len + 1, (n_reads,), generator=g, device=dev, dtype=torch.int32)
        reads = torch.where(cols[None, :] < lengths[:, None], reads, torch.zeros_like(reads))
    return reads, lengths


def make_queries(genome: torch.Tensor, n_queries: int, k: int, seed: int, frac_present: float = 0.9) -> torch.Tensor:
    """[Q, k] ASCII uint8 k-mers: `frac_present` cut from `genome`, the rest uniform random (section 8(d))."""
    dev = genome.device
    g = _gen(seed, dev)
    starts = torch.randint(0, genome.numel() - k + 1, (n_queries,), generator=g, device=dev)
    cols = torch.arange(k, device=dev)
    q = genome[starts[:, None] + cols[None, :]]
    rnd = torch.randint(1, 5, (n_queries, k), generator=g, device=dev, dtype=torch.uint8)
    use_rnd = torch.rand(n_queries, generator=g, device=dev) >= frac_present
    q = torch.where(use_rnd[:, None], rnd, q)
    lut = torch.from_numpy(ASCII.copy()).to(dev)
    return lut[q.long()]


def codes_to_ascii(codes: torch.Tensor) -> np.ndarray:
    return ASCII[codes.cpu().numpy()]


def reads_to_strings(reads: torch.Tensor, lengths: torch.Tensor) -> list[str]:
    a = codes_to_ascii(reads)
    L = lengths.cpu().numpy()
    return [a[i, : L[i]].tobytes().decode() for i in range(a.shape[0])]


def strings_to_reads(strs: list[str], device="cpu") -> tuple[torch.Tensor, torch.Tensor]:
    lut = np.zeros(256, dtype=np.uint8)
    for i, c in enumerate(b"$ACGT"):
        lut[c] = i
    L = max((len(s) for s in strs), default=0)
    out = np.zeros((len(strs), L), dtype=np.uint8)
    for i, s in enumerate(strs):
        out[i, : len(s)] = lut[np.frombuffer(s.encode(), dtype=np.uint8)]
    lens = np.array([len(s) for s in strs], dtype=np.int32)
    return torch.from_numpy(out).to(device), torch.from_numpy(lens).to(device)


# ----------------------------------------------------------------------------------------------
# multi-word LSD radix sort helpers
# ----------------------------------------------------------------------------------------------
def _lsd_argsort(n: int, nwords: int, key_of_word, device) -> torch.Tensor:
    """Stable argsort of n items by (word0, word1, ...) where key_of_word(w) returns int64[n] in item order."""
    perm = None
    for w in range(nwords - 1, -1, -1):
        keys = key_of_word(w)
        if perm is not None:
            keys = keys[perm]
        order = torch.sort(keys, stable=True).indices
        del keys
        perm = order if perm is None else perm[order]
        del order
    if perm is None:
        perm = torch.arange(n, device=device)
    return perm


def dedupe_rlo(reads: torch.Tensor, lengths: torch.Tensor):
    """Unique reads in RLO order (sorted by reversed string, shorter-is-smaller on a tie of the common
    part, as std::string::operator< does).  Returns (reads, lengths, first_index, inverse) where
    inverse[i] = position of input read i in the output."""
    n, L = reads.shape
    dev = reads.device
    if n == 0:
        return reads, lengths, torch.zeros(0, dtype=torch.long, device=dev), torch.zeros(0, dtype=torch.long, device=dev)
    cols = torch.arange(L, device=dev)
    # reversed[i, j] = reads[i, len-1-j]  (0 beyond the read)
    src = lengths[:, None].long() - 1 - cols[None, :]
    rev = torch.where(src >= 0, torch.gather(reads, 1, src.clamp(min=0)), torch.zeros_like(reads))
    del src
    nwords = (L + CHARS_PER_WORD - 1) // CHARS_PER_WORD

    def key(w):
        k = torch.zeros(n, dtype=torch.int64, device=dev)
        for j in range(w * CHARS_PER_WORD, (w + 1) * CHARS_PER_WORD):
            k <<= 3
            if j < L:
                k |= rev[:, j].long()
        return k

    perm = _lsd_argsort(n, nwords, key, dev)
    srt = rev[perm]
    del rev
    same = torch.zeros(n, dtype=torch.bool, device=dev)
    if n > 1:
        same[1:] = (srt[1:] == srt[:-1]).all(dim=1)
    del srt
    group = torch.cumsum((~same).long(), 0) - 1  # output slot of each sorted item
    keep = perm[~same]
    inverse = torch.empty(n, dtype=torch.long, device=dev)
    inverse[perm] = group
    return reads[keep], lengths[keep], keep, inverse


# ----------------------------------------------------------------------------------------------
# BWT of a read collection
# ----------------------------------------------------------------------------------------------
@dataclass
class BuiltBWT:
    runs: torch.Tensor        # uint8 run bytes (sym<<5 | len), SGA binary RLE payload
    num_strings: int
    num_symbols: int
    sai: torch.Tensor         # int64 [num_strings]: lexicographic rank -> input-order read index
    bwt: torch.Tensor | None  # uint8 [num_symbols] plain BWT codes (kept only when keep_plain=True)


def build_bwt(reads: torch.Tensor, lengths: torch.Tensor, keep_plain: bool = False) -> BuiltBWT:
    """BWT of reads (input order = read index), `$_i < $_j <=> i < j`, by LSD radix sort of all suffixes."""
    U, L = reads.shape
    dev = reads.device
    cols = torch.arange(L + 1, device=dev)
    valid = cols[None, :] <= lengths[:, None]                      # col == len is the '$'
    padded = torch.cat([reads, torch.zeros((U, 1), dtype=torch.uint8, device=dev)], dim=1)
    padded = torch.where(cols[None, :] < lengths[:, None], padded, torch.zeros_like(padded))
    T = padded[valid]                                               # text, read by read, '$' = 0
    dist = (lengths[:, None].to(torch.int16) - cols[None, :].to(torch.int16))[valid]  # distance to own '$'
    is_start = (cols[None, :] == 0).expand(U, L + 1)[valid]
    del padded, valid
    n = T.numel()
    maxlen = int(lengths.max().item()) if U else 0
    nwords = (maxlen + 1 + CHARS_PER_WORD - 1) // CHARS_PER_WORD
    Tpad = torch.cat([T, torch.zeros(nwords * CHARS_PER_WORD, dtype=torch.uint8, device=dev)])

    def key(w):
        k = torch.zeros(n, dtype=torch.int64, device=dev)
        for off in range(w * CHARS_PER_WORD, (w + 1) * CHARS_PER_WORD):
            k <<= 3
            k |= torch.where(dist > off, Tpad[off: off + n], torch.zeros((), dtype=torch.uint8, device=dev)).long()
        return k

    perm = _lsd_argsort(n, nwords, key, dev)                        # row -> text position
    del Tpad, dist
    prev = torch.cat([torch.zeros(1, dtype=torch.uint8, device=dev), T[:-1]])  # T[p-1]; '$' before a read start
    bwt = prev[perm]
    del prev
    # lexicographic rank -> read index: rows whose suffix is a whole read, in row order
    read_of_pos = torch.cumsum(is_start.long(), 0) - 1
    start_rows = is_start[perm]
    sai = read_of_pos[perm[start_rows]]
    del perm, read_of_pos, start_rows, is_start
    runs = rle_encode(bwt)
    return BuiltBWT(runs=runs, num_strings=U, num_symbols=n, sai=sai, bwt=bwt if keep_plain else None)


def rle_encode(bwt: torch.Tensor) -> torch.Tensor:
    """Plain BWT codes -> SGA RLUnit bytes; runs longer than 31 are split 31,31,...,rest
    (BWTWriterBinary.cpp:50-71 merges greedily until a unit is full)."""
    n = bwt.numel()
    dev = bwt.device
    if n == 0:
        return torch.zeros(0, dtype=torch.uint8, device=dev)
    change = torch.ones(n, dtype=torch.bool, device=dev)
    change[1:] = bwt[1:] != bwt[:-1]
    starts = torch.nonzero(change).flatten()
    del change
    sym = bwt[starts]
    ends = torch.cat([starts[1:], torch.tensor([n], device=dev)])
    rl = ends - starts
    del starts, ends
    units = (rl + 30) // 31
    total = int(units.sum().item())
    run_id = torch.repeat_interleave(torch.arange(rl.numel(), device=dev), units, output_size=total)
    first_unit = torch.cumsum(units, 0) - units
    k = torch.arange(total, device=dev) - first_unit[run_id]        # unit number within its run
    is_last = k == (units[run_id] - 1)
    ulen = torch.where(is_last, rl[run_id] - 31 * (units[run_id] - 1), torch.full_like(k, 31))
    return ((sym[run_id].long() << 5) | ulen).to(torch.uint8)


def rle_decode(runs: torch.Tensor) -> torch.Tensor:
    r = runs.long()
    return torch.repeat_interleave((r >> 5).to(torch.uint8), r & 31)


def bwt_file_bytes(runs: torch.Tensor, num_strings: int, num_symbols: int) -> bytes:
    """rlebwt_reader.cpp:27-48 / BWTWriterBinary.cpp:28-46: u16 magic, u64 x3, i32 flag(=0), run bytes."""
    hdr = struct.pack("<HQQQi", BWT_MAGIC, num_strings, num_symbols, runs.numel(), 0)
    assert len(hdr) == BWT_HEADER_BYTES
    return hdr + runs.cpu().numpy().tobytes()


def write_bwt(path: str, runs: torch.Tensor, num_strings: int, num_symbols: int) -> None:
    with open(path, "wb") as f:
        f.write(bwt_file_bytes(runs, num_strings, num_symbols))


def read_bwt(path: str):
    with open(path, "rb") as f:
        hdr = f.read(BWT_HEADER_BYTES)
        magic, ns, nsym, nruns, _flag = struct.unpack("<HQQQi", hdr)
        if magic != BWT_MAGIC:
            raise ValueError("BWT file is not properly formatted")
        runs = np.frombuffer(f.read(nruns), dtype=np.uint8)
    return runs, ns, nsym


def q2_safe(runs: torch.Tensor, num_symbols: int) -> bool:
    """the reference pushes an m_vSum hint only at the *start* of a run whose preceding
    total has reached the next 65,536 boundary (rlebwt.cpp:103-106), so vsum_size = 1 + (start of the
    last run >> 16); a lookup at index n reads m_vSum[n >> 16] and is in bounds iff the last run starts
    in the same 65,536-window that n falls in."""
    if runs.numel() == 0:
        return False
    last_len = int(runs[-1].item()) & 31
    return (num_symbols >> 16) == ((num_symbols - last_len) >> 16)


# ----------------------------------------------------------------------------------------------
# named configurations
# ----------------------------------------------------------------------------------------------
@dataclass
class Workload:
    name: str
    haplotypes: list[torch.Tensor]
    reads: torch.Tensor
    lengths: torch.Tensor
    built: BuiltBWT
    queries: torch.Tensor  # [Q, k] ASCII


def config2(device="cpu", genome_len: int = 10_000_000, n_reads: int = 50_000_000, read_len: int = 100,
            n_queries: int = 1_000_000, k: int = 31, keep_plain: bool = False) -> Workload:
    """BASELINE.json configs[1]: 50M x 100bp reads from a random 10 Mbp diploid genome (SNPs ~1/kb),
    de-duplicated, one RLBWT; 1M 31-mers, 90% cut from haplotype 1. Seeds 2 (data) and 3 (queries)."""
    h1 = random_genome(genome_len, seed=2, device=device)
    h2 = mutate_snps(h1, 1e-3, seed=20)
    reads, lens = sample_reads([h1, h2], n_reads, read_len, seed=21)
    reads, lens, _, _ = dedupe_rlo(reads, lens)
    built = build_bwt(reads, lens, keep_plain=keep_plain)
    queries = make_queries(h1, n_queries, k, seed=3)
    return Workload(f"config2:{n_reads}x{read_len}bp/{genome_len}bp", [h1, h2], reads, lens, built, queries)


# ----------------------------------------------------------------------------------------------
# population of diploid samples, reads partitioned into shards (BASELINE.json configs[2])
# ----------------------------------------------------------------------------------------------
PATTERN_BITS = 16   # sites a read window may span (population site density 1/300: a 100-bp window holds ~0.3)


@dataclass
class Population:
    """A reference genome, a shared list of variant sites and, per haplotype, which sites carry the alternative
    allele.  Sample i owns haplotypes 2i and 2i+1."""
    genome: torch.Tensor      # uint8 [G], codes 1..4
    site_pos: torch.Tensor    # int64 [M], ascending
    site_alt: torch.Tensor    # uint8 [M], alternative base (differs from the genome's)
    hap_alt: torch.Tensor     # bool [2 * n_samples, M]
    site_cum: torch.Tensor    # int32 [G + 1]: sites at positions < p
    n_samples: int

    @property
    def n_haplotypes(self) -> int:
        return 2 * self.n_samples


def make_population(genome_len: int, n_samples: int, seed: int = 4, site_rate: float = 1.0 / 300, snp_rate: float = 1e-3,
                    device="cpu") -> Population:
    """Random genome; a population site list (each position with probability site_rate); every haplotype carries the
    alternative allele at a site with probability snp_rate / site_rate, i.e. ~snp_rate substitutions per base."""
    g = _gen(seed, device)
    genome = torch.randint(1, 5, (genome_len,), generator=g, device=device, dtype=torch.uint8)
    is_site = torch.rand(genome_len, generator=g, device=device) < site_rate
    site_pos = torch.nonzero(is_site).flatten()
    m = site_pos.numel()
    shift = torch.randint(1, 4, (m,), generator=g, device=device, dtype=torch.uint8)
    site_alt = (genome[site_pos] - 1 + shift) % 4 + 1
    hap_alt = torch.rand((2 * n_samples, m), generator=g, device=device) < (snp_rate / site_rate)
    site_cum = torch.zeros(genome_len + 1, dtype=torch.int32, device=device)
    site_cum[1:] = torch.cumsum(is_site.to(torch.int32), 0)
    return Population(genome, site_pos, site_alt, hap_alt, site_cum, n_samples)


def _read_keys(pop: Population, start: torch.Tensor, hap: torch.Tensor, read_len: int):
    """Per read (start, haplotype): its shard (reversed last 3 bases as a base-4 number: the 64 keys of
    demo/permutations-3.txt in that file's order, service.cpp:820) and the pattern of alternative alleles over the
    sites inside its window -- (start, pattern) identifies the read's string."""
    m = pop.site_pos.numel()
    a = pop.site_cum[start].long()
    b = pop.site_cum[start + read_len].long()
    nsites = b - a
    if nsites.numel() and int(nsites.max()) > PATTERN_BITS:
        raise RuntimeError("a read window spans more variant sites than PATTERN_BITS")
    flat = pop.hap_alt.reshape(-1)
    pattern = torch.zeros_like(start)
    for j in range(int(nsites.max()) if nsites.numel() else 0):
        idx = (a + j).clamp(max=max(m - 1, 0))
        bit = flat[hap.long() * m + idx] & (a + j < b)
        pattern |= bit.long() << j
    # last three bases under the haplotype
    shard = torch.zeros_like(start)
    for j in range(3):
        p = start + read_len - 1 - j
        sidx = pop.site_cum[p].long()
        at_site = pop.site_cum[p + 1].long() > sidx
        sidx_c = sidx.clamp(max=max(m - 1, 0))
        alt = at_site & flat[hap.long() * m + sidx_c] if m else at_site
        base = torch.where(alt, pop.site_alt[sidx_c] if m else pop.genome[p], pop.genome[p]).long()
        shard = shard * 4 + (base - 1)
    return shard, pattern


def population_read_keys(pop: Population, coverage: float, read_len: int, seed: int, shards: list[int] | None = None,
                         n_shard_keys: int = 64):
    """All reads of the population as sorted 63-bit keys  shard << 57 | start << 26 | pattern << 10 | haplotype
    (forward strand, error free, uniform start, coverage x per diploid sample split evenly over its two haplotypes).
    Only the reads of `shards` are kept (None = all); the per-shard raw read counts of ALL shards are returned too
    (the weights for placing shards on GPUs).  haplotype < 1024, start < 2^31."""
    dev = pop.genome.device
    G = pop.genome.numel()
    per_sample = int(coverage * G / read_len)
    assert pop.n_haplotypes <= 1024 and G < (1 << 31) and n_shard_keys <= 64
    keep = None
    if shards is not None:
        keep = torch.zeros(n_shard_keys, dtype=torch.bool, device=dev)
        keep[torch.tensor(sorted(shards), device=dev, dtype=torch.long)] = True
    hist = torch.zeros(n_shard_keys, dtype=torch.int64, device=dev)
    parts = []
    for s in range(pop.n_samples):
        g = _gen(seed * 1_000_003 + s, dev)
        start = torch.randint(0, G - read_len + 1, (per_sample,), generator=g, device=dev)
        hap = torch.randint(0, 2, (per_sample,), generator=g, device=dev) + 2 * s
        shard, pattern = _read_keys(pop, start, hap, read_len)
        hist += torch.bincount(shard, minlength=n_shard_keys)
        key = (shard << 57) | (start << 26) | (pattern << 10) | hap
        if keep is not None:
            key = key[keep[shard]]
        parts.append(key)
        del start, hap, shard, pattern
    keys = torch.cat(parts) if parts else torch.zeros(0, dtype=torch.int64, device=dev)
    del parts
    keys = torch.sort(keys).values
    return keys, hist


@dataclass
class ShardData:
    shard: int                 # 0..63: index into permutation_keys(3)
    reads: torch.Tensor        # uint8 [U, L] unique reads in RLO order (= builder input order = read index)
    built: BuiltBWT
    samp_off: torch.Tensor     # int64 [U + 1]
    samp_codes: torch.Tensor   # uint8 [total]: code of sample i = 34 + i (create_samples_hash.cpp: 1-based + 33), in
                               #   sample order, one per occurrence of the read (rlosort...cpp:47-53, merge...cpp:45-47)
    raw_reads: int
    dropped_reads: int = 0     # unique reads removed to keep the reference's m_vSum lookup in bounds (Q2)


def build_population_shard(pop: Population, keys: torch.Tensor, shard: int, read_len: int, keep_plain: bool = False) -> ShardData:
    """One shard's slice of the sorted keys -> unique reads (RLO order), their BWT + .sai, and the sample table."""
    dev = pop.genome.device
    bounds = torch.tensor([shard << 57, ((shard + 1) << 57) - 1], device=dev, dtype=torch.int64)
    lo = int(torch.searchsorted(keys, bounds[:1])[0])
    hi = int(torch.searchsorted(keys, bounds[1:], right=True)[0])
    k = keys[lo:hi]
    n_raw = k.numel()
    cand_key = k >> 10                                   # (shard, start, pattern): one candidate read string each
    first = torch.ones(n_raw, dtype=torch.bool, device=dev)
    if n_raw > 1:
        first[1:] = cand_key[1:] != cand_key[:-1]
    cand_of = torch.cumsum(first.long(), 0) - 1          # raw read -> candidate
    ck = cand_key[first]
    start = (ck >> 16) & 0x7FFFFFFF
    pattern = ck & 0xFFFF
    U0 = ck.numel()
    cols = torch.arange(read_len, device=dev)
    reads = pop.genome[start[:, None] + cols[None, :]]
    a = pop.site_cum[start].long()
    m = pop.site_pos.numel()
    for j in range(PATTERN_BITS):
        sel = torch.nonzero((pattern >> j) & 1).flatten()
        if sel.numel() == 0:
            continue
        sidx = (a[sel] + j).clamp(max=max(m - 1, 0))
        reads[sel, pop.site_pos[sidx] - start[sel]] = pop.site_alt[sidx]
    lens = torch.full((U0,), read_len, dtype=torch.int32, device=dev)
    ureads, ulens, _, inverse = dedupe_rlo(reads, lens)  # exact: two windows with the same string merge here
    del reads
    built = build_bwt(ureads, ulens, keep_plain=keep_plain)
    dropped = 0
    while not q2_safe(built.runs, built.num_symbols) and ureads.shape[0] > 1:
        # the reference would read m_vSum out of bounds on this BWT (~31/65536 of all BWTs): drop the
        # shard's last read (RLO order) until it would not -- deterministic, and the generator's only other choice
        # would be to re-seed the whole population
        dropped += 1
        ureads, ulens = ureads[:-1], ulens[:-1]
        built = build_bwt(ureads, ulens, keep_plain=keep_plain)
    U = ureads.shape[0]
    if dropped:
        keep_raw = inverse[cand_of] < U
        k, cand_of = k[keep_raw], cand_of[keep_raw]
    # sample table: per unique read, the codes of the samples it was seen in, sample order, duplicates kept
    rid = inverse[cand_of]
    sample = (k & 1023) >> 1
    order = torch.sort(rid * 1024 + sample).values
    samp_codes = (order % 1024 + 34).to(torch.uint8)
    samp_off = torch.zeros(U + 1, dtype=torch.int64, device=dev)
    samp_off[1:] = torch.cumsum(torch.bincount(order // 1024, minlength=U), 0)
    return ShardData(shard, ureads, built, samp_off, samp_codes, n_raw, dropped)


def population_queries(pop: Population, n_queries: int, k: int, seed: int, frac_present: float = 0.9) -> torch.Tensor:
    """[Q, k] ASCII: frac_present cut from the reference genome (haplotypes differ from it at ~1/kb), the rest random."""
    return make_queries(pop.genome, n_queries, k, seed, frac_present)


def site_probe_kmers(pop: Population, n_sites: int, k: int, seed: int):
    """configs[4]: for n_sites sites of the population list, the reference-allele and the alternative-allele k-mer
    centred on the site.  Returns (ref [n, k] ASCII, alt [n, k] ASCII, site numbers [n])."""
    dev = pop.genome.device
    g = _gen(seed, dev)
    G = pop.genome.numel()
    half = k // 2
    ok = torch.nonzero((pop.site_pos >= half) & (pop.site_pos + (k - half) <= G)).flatten()
    pick = ok[torch.randint(0, ok.numel(), (n_sites,), generator=g, device=dev)]
    pos = pop.site_pos[pick]
    cols = torch.arange(k, device=dev)
    ref = pop.genome[(pos - half)[:, None] + cols[None, :]]
    alt = ref.clone()
    alt[:, half] = pop.site_alt[pick]
    lut = torch.from_numpy(ASCII.copy()).to(dev)
    return lut[ref.long()], lut[alt.long()], pick


# ------------------------------------------------------------------------------------------------
# run-level collections (BASELINE.json configs[3]): BWTs far too large to derive from reads here are drawn
# directly as SGA run bytes, run symbol and run length following a table measured on a real shard
# ------------------------------------------------------------------------------------------------
def run_table(runs: torch.Tensor) -> torch.Tensor:
    """[5, 32] float64: how many runs of each symbol ($ACGT) have each length (1..31) in the run bytes `runs`."""
    r = runs.to(torch.int64)
    return torch.bincount((r >> 5) * 32 + (r & 31), minlength=160)[:160].reshape(5, 32).to(torch.float64)


def runlevel_shard(table: torch.Tensor, num_symbols: int, seed: int, device="cuda", chunk: int = 1 << 27):
    """Run bytes of a BWT of about `num_symbols` symbols whose (symbol, length) pairs are drawn independently from
    `table` (run_table of a real shard).  Returns (runs uint8 [R] on `device`, num_strings, num_symbols): the number of
    strings is the number of '$' drawn, as in a real BWT.  Not the BWT of any read set -- LF walks still end at a '$'
    (geometrically distributed lengths) -- which is all the count and locate kernels need (Config 4)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    p = (table / table.sum()).reshape(-1).to(device)
    cdf = torch.cumsum(p, 0).to(torch.float32)
    mean_len = float((table * torch.arange(32, dtype=torch.float64)).sum() / table.sum())
    total = int(num_symbols / mean_len)
    out = torch.empty(total, dtype=torch.uint8, device=device)
    nsym = nstr = 0
    for a in range(0, total, chunk):
        n = min(chunk, total - a)
        u = torch.rand(n, generator=g, device=device)
        cell = torch.searchsorted(cdf, u).clamp_(max=159)          # index into the [5, 32] table
        ln = cell & 31
        bad = ln == 0                                              # (a zero-length cell has probability 0; rounding)
        cell = torch.where(bad, torch.full_like(cell, 1 * 32 + 1), cell)
        ln = cell & 31
        out[a:a + n] = cell.to(torch.uint8)                        # sym << 5 | len == cell
        nsym += int(ln.sum())
        nstr += int(ln[(cell >> 5) == 0].sum())
        del u, cell, ln, bad
    return out, nstr, nsym
