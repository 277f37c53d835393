"""Shared test data: seeded synthetic read sets -> SGA-format run-length BWTs (datagen), the oracle
handles over them, and the host replay of the device code (tests/emul).  TEST INFRASTRUCTURE."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

import oracle
from readserver_b200 import datagen as dg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMUL_DIR = os.path.join(ROOT, "tests", "emul")
EMUL_SO = os.path.join(EMUL_DIR, "librs_emul.so")
CSRC = os.path.join(ROOT, "readserver_b200", "csrc")
GOLDEN = os.path.join(ROOT, "tests", "golden")

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")


@dataclass
class Case:
    name: str
    runs: np.ndarray
    num_strings: int
    num_symbols: int
    sai: np.ndarray
    genome: object
    reads: list | None = None

    def queries(self, n: int, k: int = 31, seed: int = 5) -> np.ndarray:
        return dg.make_queries(self.genome, n, k, seed).numpy()


def make_case(name: str, genome_len: int, n_reads: int, seed: int, min_len: int | None = 73, read_len: int = 100,
              keep_reads: bool = False) -> Case:
    """Diploid genome (SNPs 1/kb), forward-strand error-free reads, de-duplicated, RLO order (the
    reference ETL's conventions); re-seeded until the index is Q2-safe."""
    for attempt in range(8):
        s = seed + 1000 * attempt
        g = dg.random_genome(genome_len, s)
        h2 = dg.mutate_snps(g, 1e-3, s + 1)
        reads, lens = dg.sample_reads([g, h2], n_reads, read_len, s + 2, min_len=min_len)
        reads, lens, _, _ = dg.dedupe_rlo(reads, lens)
        b = dg.build_bwt(reads, lens)
        if dg.q2_safe(b.runs, b.num_symbols):
            return Case(name, b.runs.numpy().copy(), b.num_strings, b.num_symbols, b.sai.numpy().astype(np.uint64), g,
                        dg.reads_to_strings(reads, lens) if keep_reads else None)
    raise RuntimeError("could not generate a Q2-safe BWT")


class Datasets:
    """Lazily built, cached per session."""

    SPECS = {
        # name: (genome_len, n_reads, seed, min_len)      reference tree depth / tail region
        "tiny": (500, 30, 11, 73),            # depth 1, one m_vSum window: everything is 'tail'
        "small": (20000, 6000, 1, 73),        # depth 2, tail region of ~55k rows
        "clean": (24000, 6800, 3, 73),        # depth 2, EMPTY tail region: the reference is exact everywhere
        "mid": (300000, 60000, 21, 73),       # depth 2, 80 windows
        "deep": (1000000, 80000, 31, 100),    # depth 3 (> 2^20 runs), empty tail region
    }

    def __init__(self):
        self._cases = {}
        self._port = {}

    def case(self, name: str) -> Case:
        if name not in self._cases:
            gl, nr, seed, ml = self.SPECS[name]
            self._cases[name] = make_case(name, gl, nr, seed, ml, keep_reads=(name in ("tiny", "small", "clean")))
        return self._cases[name]

    def port(self, name: str) -> oracle.PortIndex:
        if name not in self._port:
            c = self.case(name)
            self._port[name] = oracle.PortIndex(runs=c.runs, num_strings=c.num_strings, num_symbols=c.num_symbols)
        return self._port[name]


def write_case(case: Case, path: str) -> str:
    import torch
    dg.write_bwt(path, torch.from_numpy(case.runs), case.num_strings, case.num_symbols)
    return path


# ------------------------------------------------------------------------------------------------
# host replay of the device code (tests/emul/rs_emul.cpp over the real builder + rs_ops.h)
# ------------------------------------------------------------------------------------------------
def build_emul() -> str:
    srcs = [os.path.join(EMUL_DIR, "rs_emul.cpp"), os.path.join(CSRC, "rs_build.cpp")]
    deps = srcs + [os.path.join(CSRC, h) for h in ("rs_build.h", "rs_layout.h", "rs_swar.h", "rs_ops.h", "rs_gt_ops.h")]
    if os.path.exists(EMUL_SO) and all(os.path.getmtime(d) <= os.path.getmtime(EMUL_SO) for d in deps):
        return EMUL_SO
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-Wno-unknown-pragmas", "-I", CSRC,
                    "-o", EMUL_SO] + srcs, check=True)
    return EMUL_SO


class Emul:
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(build_emul())
            L.emul_open.restype = C.c_void_p
            L.emul_open.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_int)]
            L.emul_open_file.restype = C.c_void_p
            L.emul_open_file.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
            L.emul_close.argtypes = [C.c_void_p]
            L.emul_set_tail.argtypes = [C.c_void_p, C.c_int]
            L.emul_info.argtypes = [C.c_void_p, _u64p]
            L.emul_occ_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_uint64, _u64p]
            L.emul_char_batch.argtypes = [C.c_void_p, _u64p, C.c_uint64, _u8p]
            L.emul_occ_at_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_uint64, _u64p]
            L.emul_find_interval_batch.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_uint64, _u64p, _u64p]
            L.emul_prefix_walk_batch.argtypes = [C.c_void_p, _u64p, C.c_uint64, C.c_uint32, _u64p, _u32p]
            L.emul_postfix_walk_batch.argtypes = [C.c_void_p, _u64p, C.c_uint64, C.c_uint32, _u64p, _u32p]
            L.emul_parse_etl.restype = C.c_int
            L.emul_parse_etl.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
            L.emul_gt_seeds.argtypes = [C.c_void_p, _u8p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                        C.c_uint64, _u64p]
            L.emul_blocks.restype = C.c_void_p
            L.emul_blocks.argtypes = [C.c_void_p]
            cls._lib = L
        return cls._lib

    def __init__(self, case: Case | None = None, path: str | None = None):
        L = self.lib()
        st = C.c_int()
        if path is not None:
            self._h = L.emul_open_file(path.encode(), C.byref(st))
        else:
            self._h = L.emul_open(case.runs, case.runs.size, case.num_strings, case.num_symbols, C.byref(st))
        self.status = st.value
        if self._h:
            info = np.zeros(32, dtype=np.uint64)
            L.emul_info(self._h, info)
            self.n, self.nblocks, self.ndir, self.clamp_c, self.vlast, self.depth, self.q2_safe = (int(x) for x in info[:7])
            self.pc = [int(x) for x in info[7:12]]
            self.tot = [int(x) for x in info[12:17]]
            self.init_occ = [int(x) for x in info[17:22]]
            self.clamp_char = int(info[22])

    def __del__(self):
        if getattr(self, "_h", None):
            self.lib().emul_close(self._h)
            self._h = None

    def set_tail(self, on: bool):
        self.lib().emul_set_tail(self._h, int(on))

    def occ(self, syms, idx):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        syms = np.ascontiguousarray(syms, dtype=np.uint8)
        out = np.empty(idx.size, dtype=np.uint64)
        self.lib().emul_occ_batch(self._h, syms, idx, idx.size, out)
        return out

    def char(self, idx):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.empty(idx.size, dtype=np.uint8)
        self.lib().emul_char_batch(self._h, idx, idx.size, out)
        return out

    def occ_at(self, syms, cnt):
        cnt = np.ascontiguousarray(cnt, dtype=np.uint64)
        syms = np.ascontiguousarray(syms, dtype=np.uint8)
        out = np.empty(cnt.size, dtype=np.uint64)
        self.lib().emul_occ_at_batch(self._h, syms, cnt, cnt.size, out)
        return out

    def find_interval(self, kmers, k):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint8).reshape(-1)
        q = kmers.size // k
        lo = np.empty(q, dtype=np.uint64)
        up = np.empty(q, dtype=np.uint64)
        self.lib().emul_find_interval_batch(self._h, kmers, k, q, lo, up)
        return lo, up

    def gt_seeds(self, w: str, pos: int, kmers: int, skip: int, max_iv: int):
        """per tile of window w: (lo0, hi0, thr0, lo1, hi1, thr1, kind, flags), the seed logic of rs_gt_ops.h"""
        wb = np.frombuffer(w.encode(), dtype=np.uint8).copy()
        nt = (len(w) - kmers) // (skip + 1) + 1 if len(w) >= kmers and pos <= len(w) else 0
        out = np.zeros((nt, 8), dtype=np.uint64)
        if nt:
            self.lib().emul_gt_seeds(self._h, wb, wb.size, pos, nt, kmers, skip + 1, max_iv, out.reshape(-1))
        return out

    def blocks(self) -> np.ndarray:
        """the raw 64-byte blocks as a [nblocks, 64] uint8 array (a copy)"""
        p = self.lib().emul_blocks(self._h)
        buf = (C.c_uint8 * (self.nblocks * 64)).from_address(p)
        return np.frombuffer(buf, dtype=np.uint8).reshape(self.nblocks, 64).copy()

    def postfix_walk(self, rows, max_steps=1024):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        ri = np.empty(rows.size, dtype=np.uint64)
        st = np.empty(rows.size, dtype=np.uint32)
        self.lib().emul_postfix_walk_batch(self._h, rows, rows.size, max_steps, ri, st)
        return ri, st

    def prefix_walk(self, rows, max_steps=1024):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        lex = np.empty(rows.size, dtype=np.uint64)
        st = np.empty(rows.size, dtype=np.uint32)
        self.lib().emul_prefix_walk_batch(self._h, rows, rows.size, max_steps, lex, st)
        return lex, st


def write_etl_file(path: str, reads: list, off, codes, sos: int, meta: bool, n_reads_every: int = 0, seed: int = 9):
    """merged.rlosorted.reads as the reference's ETL writes it: `read\\n value\\n` per read, value = codes in
    merge order, each followed by count+33 and lane+33 bytes iff meta (service.cpp:829-843).  With n_reads_every,
    a read holding 'N' (dropped before the BWT is built, build_bwt.sh:919) is inserted every that many lines."""
    rng = np.random.default_rng(seed)
    codes = np.ascontiguousarray(codes, dtype=np.uint8).reshape(-1, sos)
    with open(path, "wb") as f:
        for i, r in enumerate(reads):
            if n_reads_every and i % n_reads_every == 0:
                f.write(b"ACGTNNACGT" + r.encode()[:20] + b"\n" + bytes(oracle.sample_code(1, sos)) + (b"\x22\x22" if meta else b"") + b"\n")
            val = b""
            for c in codes[int(off[i]):int(off[i + 1])]:
                val += bytes(c)
                if meta:
                    val += bytes([33 + int(rng.integers(0, 90)), 33 + int(rng.integers(0, 8))])
            f.write(r.encode() + b"\n" + val + b"\n")


def parse_etl(path: str, sos: int, meta: bool):
    """rs_parse_etl_samples through the test library: (status, offsets, codes)."""
    L = Emul.lib()
    nr, nc = C.c_uint64(), C.c_uint64()
    rc = L.emul_parse_etl(path.encode(), sos, int(meta), None, 0, None, 0, C.byref(nr), C.byref(nc))
    if rc != 0:
        return rc, None, None
    off = np.zeros(nr.value + 1, dtype=np.uint64)
    codes = np.zeros(max(nc.value, 1), dtype=np.uint8)
    rc = L.emul_parse_etl(path.encode(), sos, int(meta), off.ctypes.data, off.size, codes.ctypes.data, codes.size,
                          C.byref(nr), C.byref(nc))
    return rc, off, codes[: nc.value]


def positions(n: int, limit: int = 400_000, seed: int = 0) -> np.ndarray:
    """All positions of a small BWT; for a large one a random sample plus both ends (the reference's
    tail region sits at the end)."""
    if n <= limit:
        return np.arange(n, dtype=np.uint64)
    rng = np.random.default_rng(seed)
    return np.concatenate([rng.integers(0, n, limit // 2).astype(np.uint64),
                           np.arange(n - limit // 4, n, dtype=np.uint64),
                           np.arange(0, limit // 4, dtype=np.uint64)])


def brute_count(reads: list[str], kmer: str) -> int:
    """Occurrences of kmer as a substring of the reads (truth for findInterval counts)."""
    total = 0
    for r in reads:
        i = r.find(kmer)
        while i >= 0:
            total += 1
            i = r.find(kmer, i + 1)
    return total
