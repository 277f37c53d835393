"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the CPU oracle.

Two implementations behind one interface:

* ``PortIndex``  -- ``oracle/liboracle_port.so``: our plain-C restatement (``rs_oracle.c``).
* ``RefIndex``   -- ``oracle/_ref/libref_bwt.so``: the UNMODIFIED reference sources compiled by
  ``oracle/Makefile`` behind ``ref_shim.cpp`` (available when that .so has been built in the build
  container; it travels to the GPU box because ``oracle/_ref/`` is git-ignored, not gpurun-ignored).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package.  The product (``readserver_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_port.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libref_bwt.so")
REF_O0_SO = os.path.join(REF_DIR, "libref_bwt_O0.so")
REF_BCR_SO = os.path.join(REF_DIR, "libref_bcr.so")
REF_INDEX_RLEBWT = os.path.join(REF_DIR, "index_rlebwt")
REFERENCE_ROOT = "/root/reference"

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def build(ref: bool | None = None) -> None:
    """Compile the port (always) and, when /root/reference is present, the reference libs."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    if ref is None:
        ref = os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "bwt"))
    if ref:
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def have_bcr() -> bool:
    return os.path.exists(REF_BCR_SO)


def _ascii(a) -> np.ndarray:
    """bytes / str / uint8 array -> contiguous uint8 array."""
    if isinstance(a, str):
        a = a.encode()
    if isinstance(a, (bytes, bytearray)):
        return np.frombuffer(bytes(a), dtype=np.uint8).copy()
    return np.ascontiguousarray(a, dtype=np.uint8)


class _Base:
    """Common array-level API.  Subclasses provide the scalar/batch C calls."""

    def find_interval_batch(self, kmers, k: int, nthreads: int = 1):
        kmers = _ascii(kmers).reshape(-1)
        assert kmers.size % k == 0
        q = kmers.size // k
        lo = np.empty(q, dtype=np.uint64)
        up = np.empty(q, dtype=np.uint64)
        secs = self._fi(self._h, kmers, k, q, lo, up, nthreads)
        self.last_seconds = secs
        return lo, up

    def find_interval(self, w):
        w = _ascii(w)
        lo, up = self.find_interval_batch(w, len(w))
        return int(lo[0]), int(up[0])

    def locate_batch(self, rows, nthreads: int = 1, prefix: bool = True, postfix: bool = True,
                     max_steps: int = 1024):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        n = rows.size
        lex = np.empty(n, dtype=np.uint64) if prefix else None
        ps = np.empty(n, dtype=np.uint32) if prefix else None
        ri = np.empty(n, dtype=np.uint64) if postfix else None
        qs = np.empty(n, dtype=np.uint32) if postfix else None
        p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        secs = self._loc(self._h, rows, n, p(lex), p(ps), p(ri), p(qs), nthreads, max_steps)
        self.last_seconds = secs
        return lex, ps, ri, qs

    def get_occ_batch(self, syms, idx):
        syms = _ascii(syms)
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.empty(idx.size, dtype=np.uint64)
        self._occb(self._h, syms, idx, idx.size, out)
        return out

    def get_char_batch(self, idx):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.empty(idx.size, dtype=np.uint8)
        self._chb(self._h, idx, idx.size, out)
        return out

    def get_occ_at_batch(self, syms, cnt):
        syms = _ascii(syms)
        cnt = np.ascontiguousarray(cnt, dtype=np.uint64)
        out = np.empty(cnt.size, dtype=np.uint64)
        self._occatb(self._h, syms, cnt, cnt.size, out)
        return out


class PortIndex(_Base):
    kind = "port"
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            if not os.path.exists(PORT_SO):
                build(ref=False)
            L = C.CDLL(PORT_SO)
            L.rso_open.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
            L.rso_from_runs.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p)]
            L.rso_close.argtypes = [C.c_void_p]
            L.rso_write_bpi2.argtypes = [C.c_void_p, C.c_char_p]
            for f in ("rso_bwlen", "rso_num_strings", "rso_num_runs", "rso_depth", "rso_vsum_size",
                      "rso_q2_violations"):
                getattr(L, f).argtypes = [C.c_void_p]
                getattr(L, f).restype = C.c_uint64
            L.rso_get_pc.argtypes = [C.c_void_p, C.c_char]
            L.rso_get_pc.restype = C.c_uint64
            L.rso_get_occ.argtypes = [C.c_void_p, C.c_char, C.c_uint64]
            L.rso_get_occ.restype = C.c_uint64
            L.rso_get_occ_at.argtypes = [C.c_void_p, C.c_char, C.c_uint64]
            L.rso_get_occ_at.restype = C.c_uint64
            L.rso_get_char.argtypes = [C.c_void_p, C.c_uint64]
            L.rso_get_char.restype = C.c_char
            L.rso_get_f.argtypes = [C.c_void_p, C.c_uint64]
            L.rso_get_f.restype = C.c_char
            L.rso_extract_prefix.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t, C.POINTER(C.c_uint64), C.c_size_t]
            L.rso_extract_prefix.restype = C.c_size_t
            L.rso_extract_postfix.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t, C.POINTER(C.c_uint64), C.c_size_t]
            L.rso_extract_postfix.restype = C.c_size_t
            L.rso_find_interval_batch.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_size_t, _u64p, _u64p, C.c_int]
            L.rso_find_interval_batch.restype = C.c_double
            L.rso_locate_batch.argtypes = [C.c_void_p, _u64p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int, C.c_uint32]
            L.rso_locate_batch.restype = C.c_double
            L.rso_get_occ_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_size_t, _u64p]
            L.rso_get_occ_batch.restype = C.c_double
            L.rso_get_char_batch.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u8p]
            L.rso_get_char_batch.restype = C.c_double
            L.rso_get_occ_at_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_size_t, _u64p]
            L.rso_get_occ_at_batch.restype = C.c_double
            L.rso_decode_sample_value.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                  C.c_void_p, C.c_size_t]
            L.rso_decode_sample_value.restype = C.c_size_t
            L.rso_sample_code.argtypes = [C.c_uint64, C.c_int, _u8p]
            cls._lib = L
        return cls._lib

    def __init__(self, bwt_path: str | None = None, *, runs=None, num_strings=0, num_symbols=0,
                 use_bpi2: bool = False):
        L = self.lib()
        h = C.c_void_p()
        if bwt_path is not None:
            rc = L.rso_open(bwt_path.encode(), int(use_bpi2), C.byref(h))
        else:
            runs = np.ascontiguousarray(runs, dtype=np.uint8)
            rc = L.rso_from_runs(runs, runs.size, num_strings, num_symbols, C.byref(h))
        if rc != 0:
            raise RuntimeError(f"rso_open failed with status {rc}")
        self._h = h
        self._fi, self._loc = L.rso_find_interval_batch, L.rso_locate_batch
        self._occb, self._chb, self._occatb = L.rso_get_occ_batch, L.rso_get_char_batch, L.rso_get_occ_at_batch
        self.last_seconds = 0.0

    def close(self):
        if getattr(self, "_h", None):
            self.lib().rso_close(self._h)
            self._h = None

    __del__ = close

    @property
    def bwlen(self): return int(self.lib().rso_bwlen(self._h))
    @property
    def num_strings(self): return int(self.lib().rso_num_strings(self._h))
    @property
    def num_runs(self): return int(self.lib().rso_num_runs(self._h))
    @property
    def depth(self): return int(self.lib().rso_depth(self._h))
    @property
    def vsum_size(self): return int(self.lib().rso_vsum_size(self._h))
    @property
    def q2_violations(self): return int(self.lib().rso_q2_violations(self._h))

    def write_bpi2(self, path: str):
        rc = self.lib().rso_write_bpi2(self._h, path.encode())
        if rc:
            raise RuntimeError("rso_write_bpi2 failed")

    def get_pc(self, b): return int(self.lib().rso_get_pc(self._h, _c(b)))
    def get_occ(self, b, i): return int(self.lib().rso_get_occ(self._h, _c(b), i & 0xFFFFFFFFFFFFFFFF))
    def get_occ_at(self, b, c): return int(self.lib().rso_get_occ_at(self._h, _c(b), c))
    def get_char(self, i): return self.lib().rso_get_char(self._h, i).decode("latin1")
    def get_f(self, i): return self.lib().rso_get_f(self._h, i).decode("latin1")

    def extract_prefix(self, row: int, cap: int = 4096):
        buf = C.create_string_buffer(cap)
        term = C.c_uint64()
        n = self.lib().rso_extract_prefix(self._h, row, buf, cap, C.byref(term), cap - 2)
        return buf.value.decode("latin1"), int(term.value), int(n)

    def extract_postfix(self, row: int, cap: int = 4096):
        buf = C.create_string_buffer(cap)
        term = C.c_uint64()
        n = self.lib().rso_extract_postfix(self._h, row, buf, cap, C.byref(term), cap - 2)
        return buf.value.decode("latin1"), int(term.value), int(n)

    def extract_read(self, row: int) -> str:
        return self.extract_prefix(row)[0] + self.extract_postfix(row)[0]


def _c(b) -> bytes:
    if isinstance(b, str):
        b = b.encode("latin1")
    if isinstance(b, int):
        b = bytes([b])
    return b


class RefIndex(_Base):
    """The unmodified reference behind ref_shim.cpp."""
    kind = "reference"
    _libs: dict = {}

    @classmethod
    def lib(cls, so=REF_SO):
        if so not in cls._libs:
            if not os.path.exists(so):
                raise FileNotFoundError(f"{so} not built (make -C oracle ref needs /root/reference)")
            L = C.CDLL(so)
            L.ref_open.argtypes = [C.c_char_p]
            L.ref_open.restype = C.c_void_p
            L.ref_close.argtypes = [C.c_void_p]
            L.ref_serialise.argtypes = [C.c_void_p, C.c_char_p]
            L.ref_bwlen.argtypes = [C.c_void_p]
            L.ref_bwlen.restype = C.c_uint64
            L.ref_pc.argtypes = [C.c_void_p, C.c_char]
            L.ref_pc.restype = C.c_uint64
            L.ref_get_occ.argtypes = [C.c_void_p, C.c_char, C.c_uint64]
            L.ref_get_occ.restype = C.c_uint64
            L.ref_get_occ_at.argtypes = [C.c_void_p, C.c_char, C.c_uint64]
            L.ref_get_occ_at.restype = C.c_uint64
            L.ref_get_char.argtypes = [C.c_void_p, C.c_uint64]
            L.ref_get_char.restype = C.c_char
            L.ref_get_f.argtypes = [C.c_void_p, C.c_uint64]
            L.ref_get_f.restype = C.c_char
            L.ref_get_occ_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_size_t, _u64p]
            L.ref_get_char_batch.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u8p]
            L.ref_get_occ_at_batch.argtypes = [C.c_void_p, _u8p, _u64p, C.c_size_t, _u64p]
            L.ref_find_interval_batch.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_size_t, _u64p, _u64p, C.c_int]
            L.ref_find_interval_batch.restype = C.c_double
            L.ref_locate_batch.argtypes = [C.c_void_p, _u64p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int, C.c_uint32]
            L.ref_locate_batch.restype = C.c_double
            L.ref_extract_read.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
            L.ref_extract_read.restype = C.c_size_t
            L.ref_extract_batch_timed.argtypes = [C.c_void_p, _u64p, C.c_size_t, C.c_int, C.POINTER(C.c_uint64)]
            L.ref_extract_batch_timed.restype = C.c_double
            L.ref_query.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_size_t]
            L.ref_query.restype = C.c_size_t
            L.ref_query_exactmatch.argtypes = [C.c_void_p, C.c_char_p]
            L.ref_query_exactmatch.restype = C.c_int
            cls._libs[so] = L
        return cls._libs[so]

    def __init__(self, bwt_path: str, so: str = REF_SO):
        self._L = L = self.lib(so)
        self._h = L.ref_open(bwt_path.encode())
        self._fi, self._loc = L.ref_find_interval_batch, L.ref_locate_batch
        self._occb, self._chb, self._occatb = L.ref_get_occ_batch, L.ref_get_char_batch, L.ref_get_occ_at_batch
        self.last_seconds = 0.0

    def close(self):
        if getattr(self, "_h", None):
            self._L.ref_close(self._h)
            self._h = None

    __del__ = close

    @property
    def bwlen(self): return int(self._L.ref_bwlen(self._h))

    def serialise(self, path: str): self._L.ref_serialise(self._h, path.encode())
    def get_pc(self, b): return int(self._L.ref_pc(self._h, _c(b)))
    def get_occ(self, b, i): return int(self._L.ref_get_occ(self._h, _c(b), i & 0xFFFFFFFFFFFFFFFF))
    def get_occ_at(self, b, c): return int(self._L.ref_get_occ_at(self._h, _c(b), c))
    def get_char(self, i): return self._L.ref_get_char(self._h, i).decode("latin1")
    def get_f(self, i): return self._L.ref_get_f(self._h, i).decode("latin1")

    def extract_read(self, row: int, cap: int = 4096):
        buf = C.create_string_buffer(cap)
        pl = C.c_size_t()
        self._L.ref_extract_read(self._h, row, buf, cap, C.byref(pl))
        return buf.value.decode("latin1")

    def extract_batch_timed(self, rows, nthreads: int = 1):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        tot = C.c_uint64()
        secs = self._L.ref_extract_batch_timed(self._h, rows, rows.size, nthreads, C.byref(tot))
        return secs, int(tot.value)

    def query(self, w: str, cap: int = 1 << 24):
        buf = C.create_string_buffer(cap)
        n = self._L.ref_query(self._h, w.encode(), buf, cap)
        s = buf.value.decode("latin1")
        out = s.split("\n")[:-1] if s else []
        assert len(out) == n, "ref_query buffer too small"
        return out

    def query_exactmatch(self, w: str) -> bool:
        return bool(self._L.ref_query_exactmatch(self._h, w.encode()))


def ref_bcr_build(reads: list[str], out_bwt: str):
    """Build an SGA-format .bwt with the reference's own BWT builder (bcr.c)."""
    L = C.CDLL(REF_BCR_SO)
    L.ref_bcr_build.argtypes = [C.c_char_p, _u64p, C.c_uint64, C.c_char_p, C.c_char_p]
    L.ref_bcr_build.restype = C.c_int
    cat = "".join(reads).encode()
    off = np.zeros(len(reads) + 1, dtype=np.uint64)
    np.cumsum([len(r) for r in reads], out=off[1:])
    rc = L.ref_bcr_build(cat, off, len(reads), out_bwt.encode(), (out_bwt + ".tmp").encode())
    if rc:
        raise RuntimeError(f"ref_bcr_build failed: {rc}")


def open_best(bwt_path: str):
    """Reference if its .so is present, else the port. Used by bench.py's CPU legs."""
    if have_ref():
        return RefIndex(bwt_path)
    return PortIndex(bwt_path)


def sample_code(sample_idx: int, size_of_sample: int) -> bytes:
    out = np.zeros(size_of_sample, dtype=np.uint8)
    PortIndex.lib().rso_sample_code(sample_idx, size_of_sample, out)
    return out.tobytes()


def decode_sample_value(value: bytes, size_of_sample: int, has_other_meta_data: bool):
    v = _ascii(value)
    cap = max(1, len(value))
    codes = np.zeros(cap * size_of_sample, dtype=np.uint8)
    counts = np.zeros(cap, dtype=np.int32)
    lanes = np.zeros(cap, dtype=np.int32)
    if v.size == 0:
        return []
    n = PortIndex.lib().rso_decode_sample_value(v, v.size, size_of_sample, int(has_other_meta_data),
                                                codes.ctypes.data, counts.ctypes.data, lanes.ctypes.data, cap)
    return [(codes[i * size_of_sample:(i + 1) * size_of_sample].tobytes(), int(counts[i]), int(lanes[i]))
            for i in range(n)]


# ------------------------------------------------------------------------------------------------
# functions of the reference's query worker, its sample-code tool and its config library (oracle/_ref, Makefile `ref`)
# ------------------------------------------------------------------------------------------------
REF_SERVICE_SO = os.path.join(REF_DIR, "libref_service.so")
REF_CREATE_SAMPLES_HASH = os.path.join(REF_DIR, "create_samples_hash")
REF_CONFIG_SO = os.path.join(REF_DIR, "libref_config.so")


def have_ref_service() -> bool:
    return os.path.exists(REF_SERVICE_SO)


class RefService:
    """is_suffix_of / get_tiles / rev_comp / find_gt_reads exactly as src/service/service.cpp holds them: cut out of that
    file at build time (oracle/Makefile) and compiled behind ref_service_shim.cpp, over the reference's own RLEBWT."""
    _lib = None
    THROW = (1 << 64) - 1

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(REF_SERVICE_SO)
            sz = C.c_size_t
            L.refsvc_open.restype = C.c_void_p
            L.refsvc_open.argtypes = [C.c_char_p]
            L.refsvc_close.argtypes = [C.c_void_p]
            L.refsvc_set_limits.argtypes = [sz, sz, sz]
            L.refsvc_get_tiles.restype = sz
            L.refsvc_get_tiles.argtypes = [C.c_char_p, sz, sz, C.c_char_p, sz]
            L.refsvc_is_suffix_of.argtypes = [C.c_char_p, C.c_char_p]
            L.refsvc_rev_comp.restype = sz
            L.refsvc_rev_comp.argtypes = [C.c_char_p, C.c_char_p, sz]
            L.refsvc_find_gt_reads.restype = sz
            L.refsvc_find_gt_reads.argtypes = [C.c_void_p, C.c_char_p, sz, sz, sz, sz, C.c_char_p, sz]
            L.refsvc_gt_task_reads.restype = sz
            L.refsvc_gt_task_reads.argtypes = [C.c_void_p, C.c_char_p, sz, sz, sz, C.c_char_p, sz]
            cls._lib = L
        return cls._lib

    def __init__(self, bwt_path: str | None = None):
        self._L = self.lib()
        self._h = self._L.refsvc_open(bwt_path.encode()) if bwt_path else None

    def close(self):
        if getattr(self, "_h", None):
            self._L.refsvc_close(self._h)
            self._h = None

    __del__ = close

    def set_limits(self, max_interval_size: int = 10000, max_read_length: int = 100, min_read_length: int = 73):
        self._L.refsvc_set_limits(max_interval_size, max_read_length, min_read_length)

    def _strings(self, n, buf):
        if n == self.THROW:
            return None
        assert n != self.THROW - 1, "buffer too small"
        s = buf.value.decode("latin1")
        out = s.split("\n")[:-1] if s else []
        assert len(out) == n
        return out

    def get_tiles(self, w: str, kmer: int, skip: int = 0, cap: int = 1 << 22):
        buf = C.create_string_buffer(cap)
        return self._strings(self._L.refsvc_get_tiles(w.encode("latin1"), kmer, skip, buf, cap), buf)

    def is_suffix_of(self, s: str, w: str) -> bool:
        return bool(self._L.refsvc_is_suffix_of(s.encode(), w.encode()))

    def rev_comp(self, w: str) -> str:
        buf = C.create_string_buffer(len(w) + 8)
        self._L.refsvc_rev_comp(w.encode("latin1"), buf, len(w) + 8)
        return buf.value.decode("latin1")

    def find_gt_reads(self, w: str, index: int, kmers: int, skip: int, pos: int, cap: int = 1 << 24):
        """Sorted read strings of tile `index`; None when the reference throws (substr out of range)."""
        buf = C.create_string_buffer(cap)
        return self._strings(self._L.refsvc_find_gt_reads(self._h, w.encode("latin1"), index, kmers, skip, pos, buf, cap), buf)

    def gt_task_reads(self, w: str, kmers: int, skip: int, pos: int, cap: int = 1 << 24):
        buf = C.create_string_buffer(cap)
        return self._strings(self._L.refsvc_gt_task_reads(self._h, w.encode("latin1"), kmers, skip, pos, buf, cap), buf)


def ref_create_samples_hash(names: list[str]) -> dict:
    """name -> code bytes, by running the reference's own tool (src/util/create_samples_hash.cpp) on a name list."""
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src, dst = os.path.join(d, "names"), os.path.join(d, "hash")
        with open(src, "w") as f:
            f.write("\n".join(names) + "\n")
        subprocess.run([REF_CREATE_SAMPLES_HASH, src, dst], check=True)
        out = {}
        with open(dst, "rb") as f:
            for line in f.read().split(b"\n"):
                if line:
                    name, code = line.split(b"\t", 1)
                    out[name.decode()] = code
        return out


def ref_config(path: str, string_keys: list[str], array_keys: list[str]) -> dict:
    """The settings of a worker .cfg as the reference's libconfig 1.5 reads them (service.cpp:1407-1437)."""
    L = C.CDLL(REF_CONFIG_SO)
    L.refcfg_lookup_string.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
    L.refcfg_lookup_array.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
    buf = C.create_string_buffer(1 << 16)
    out = {}
    for k in string_keys:
        rc = L.refcfg_lookup_string(path.encode(), k.encode(), buf, len(buf))
        if rc < 0:
            raise ValueError("libconfig: parse error")
        out[k] = buf.value.decode() if rc == 1 else None
    for k in array_keys:
        n = L.refcfg_lookup_array(path.encode(), k.encode(), buf, len(buf))
        if n == -1:
            raise ValueError("libconfig: parse error")
        out[k] = buf.value.decode().split("\n")[:-1] if n >= 0 else None
    return out
