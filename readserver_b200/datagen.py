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
