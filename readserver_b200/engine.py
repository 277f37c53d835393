"""Host-side mirror of the reference's BWT interface over the C-ABI library (``librs_b200.so``).

``GpuBWT`` mirrors ``class BWT`` / ``RLEBWT`` (``/root/reference/include/bwt/bwt.h:6-15``,
``rlebwt.h:15-61``) and the free functions of ``include/bwt/query.h:18-32`` -- same names (snake
case), argument meaning and results -- in batched form.  Every call goes through the C ABI declared
in ``include/readserver_b200.h``; nothing here computes on the CPU and there is no fallback: if the
library is missing or no CUDA device is present, construction raises.

torch is used only as plumbing (device buffers and streams for the ``*_dev`` calls).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")

STATUS = {0: "OK", 1: "OPEN", 2: "MAGIC", 3: "SHORT", 4: "RUNBYTE", 5: "ARG", 6: "CUDA", 7: "TOO_LARGE", 8: "NO_TABLE", 9: "CAPACITY"}


GT_FLAG_THROW, GT_FLAG_HANG, GT_FLAG_WALK = 1, 2, 4   # rs_gt_reads_batch flags (include/readserver_b200.h)


class RsError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"readserver_b200: status {status} ({STATUS.get(status, '?')}): {msg}")
        self.status = status


class Info(C.Structure):
    _fields_ = [("bwlen", C.c_uint64), ("num_strings", C.c_uint64), ("num_runs", C.c_uint64),
                ("num_blocks", C.c_uint64), ("hbm_bytes", C.c_uint64), ("pc", C.c_uint64 * 5),
                ("total", C.c_uint64 * 5), ("ref_depth", C.c_uint64), ("ref_clamp_c", C.c_uint64),
                ("ref_vlast", C.c_uint64), ("q2_safe", C.c_int32), ("device", C.c_int32),
                ("prefix_table_k", C.c_int32), ("dense_blocks", C.c_int32)]


class ShardSetInfo(C.Structure):
    _fields_ = [("num_shards", C.c_uint32), ("sealed", C.c_uint32), ("locate_every", C.c_uint32),
                ("max_read_len", C.c_uint32), ("num_samples", C.c_uint32), ("carrier_words", C.c_uint32),
                ("index_bytes", C.c_uint64), ("table_bytes", C.c_uint64), ("bwlen", C.c_uint64),
                ("num_strings", C.c_uint64), ("num_runs", C.c_uint64), ("chunk_queries", C.c_uint64)]


_LIB = None


def lib() -> C.CDLL:
    """Loads (building if stale and nvcc is present) the C-ABI library. Raises if it cannot."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB
    if not os.path.exists(path):
        path = _build.build()
    L = C.CDLL(path)
    vp, sz, u64, i32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int
    sig = {
        "rs_index_open": (i32, [C.c_char_p, i32, C.POINTER(vp)]),
        "rs_index_from_runs": (i32, [vp, u64, u64, u64, i32, C.POINTER(vp)]),
        "rs_index_from_runs_dev": (i32, [vp, u64, u64, u64, i32, C.POINTER(vp)]),
        "rs_index_from_runs_host": (i32, [vp, u64, u64, u64, i32, C.POINTER(vp)]),
        "rs_index_save": (i32, [vp, C.c_char_p]),
        "rs_index_open_encoded": (i32, [C.c_char_p, i32, C.POINTER(vp)]),
        "rs_index_export": (i32, [vp, i32, vp, u64, C.POINTER(u64)]),
        "rs_index_set_sai_dev": (i32, [vp, vp, u64]),
        "rs_index_set_samples_dev": (i32, [vp, vp, vp, u64, i32]),
        "rs_index_set_samples_meta": (i32, [vp, vp, vp, vp, u64, i32]),
        "rs_samples_meta_batch": (i32, [vp, vp, sz, vp, u64, vp, vp]),
        "rs_shardset_create": (i32, [i32, C.POINTER(vp)]),
        "rs_shardset_add": (i32, [vp, vp, C.c_uint32]),
        "rs_shardset_seal": (i32, [vp, C.c_uint32, C.c_uint32]),
        "rs_shardset_destroy": (None, [vp]),
        "rs_shardset_get_info": (i32, [vp, C.POINTER(ShardSetInfo)]),
        "rs_shardset_enable_stats": (i32, [vp, i32]),
        "rs_shardset_read_stats": (i32, [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), i32]),
        "rs_shardset_enable_timing": (i32, [vp, i32]),
        "rs_shardset_read_timing": (i32, [vp, C.POINTER(C.c_double), C.POINTER(u64)]),
        "rs_popcomm_create": (i32, [i32, i32, i32, u64, u64, C.POINTER(vp)]),
        "rs_popcomm_export": (i32, [vp, vp]),
        "rs_popcomm_connect": (i32, [vp, vp]),
        "rs_popcomm_destroy": (None, [vp]),
        "rs_pop_query_comm_dev": (i32, [vp, vp, vp, i32, sz, i32, i32, u64, i32, i32, vp]),
        "rs_popcomm_results": (i32, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(u64)]),
        "rs_popcomm_error": (i32, [vp, C.POINTER(i32)]),
        "rs_pop_query_dev": (i32, [vp, vp, i32, sz, i32, i32, u64, vp, vp, vp, u64, vp, vp, vp]),
        "rs_pop_query": (i32, [vp, vp, i32, sz, i32, i32, u64, vp, vp, vp, u64, vp, C.POINTER(u64)]),
        "rs_pop_query_hits40": (i32, [vp, vp, i32, sz, i32, i32, u64, vp, vp, vp, u64, vp, C.POINTER(u64)]),
        "rs_index_close": (None, [vp]),
        "rs_index_get_info": (i32, [vp, C.POINTER(Info)]),
        "rs_index_set_tail_emulation": (i32, [vp, i32]),
        "rs_index_set_prefix_table": (i32, [vp, i32]),
        "rs_index_set_dense": (i32, [vp, i32]),
        "rs_index_enable_stats": (i32, [vp, i32]),
        "rs_index_read_stats": (i32, [vp, C.POINTER(u64), C.POINTER(u64), i32]),
        "rs_measure_random_access": (i32, [i32, u64, u64, C.POINTER(C.c_double)]),
        "rs_measure_random_access_granule": (i32, [i32, u64, u64, i32, C.POINTER(C.c_double)]),
        "rs_last_error": (C.c_char_p, []),
        "rs_host_alloc": (i32, [C.POINTER(vp), sz]),
        "rs_host_free": (None, [vp]),
        "rs_index_load_samples_file": (i32, [vp, C.c_char_p, i32, i32]),
        "rs_index_set_reads": (i32, [vp, vp, vp, u64]),
        "rs_index_load_reads_file": (i32, [vp, C.c_char_p]),
        "rs_reads_batch": (i32, [vp, vp, sz, C.c_uint32, vp, vp]),
        "rs_reads_batch_dev": (i32, [vp, vp, sz, C.c_uint32, vp, vp, vp]),
        "rs_tile_intervals_batch": (i32, [vp, vp, vp, sz, C.c_uint32, C.c_uint32, vp, u64, vp, vp]),
        "rs_index_build_locate_table": (i32, [vp, C.c_uint32]),
        "rs_index_build_locate_samples": (i32, [vp, C.c_uint32, C.c_uint32]),
        "rs_index_drop_locate_table": (i32, [vp]),
        "rs_gt_reads_batch": (i32, [vp, vp, vp, vp, sz, C.c_uint32, C.c_uint32, u64, C.c_uint32, u64, vp, u64, vp, vp, vp]),
        "rs_comm_create": (i32, [i32, i32, i32, u64, C.POINTER(vp)]),
        "rs_comm_export": (i32, [vp, vp]),
        "rs_comm_connect": (i32, [vp, vp]),
        "rs_comm_destroy": (None, [vp]),
        "rs_comm_begin": (i32, [vp, u64, vp]),
        "rs_count_accumulate_dev": (i32, [vp, vp, vp, i32, sz, i32, vp]),
        "rs_comm_finish": (i32, [vp, vp, C.POINTER(vp)]),
        "rs_comm_error": (i32, [vp, C.POINTER(i32)]),
        "rs_get_pc": (u64, [vp, C.c_char]),
        "rs_get_bwlen": (u64, [vp]),
        "rs_get_occ_batch": (i32, [vp, vp, vp, sz, vp]),
        "rs_get_char_batch": (i32, [vp, vp, sz, vp]),
        "rs_get_f_batch": (i32, [vp, vp, sz, vp]),
        "rs_get_occ_at_batch": (i32, [vp, vp, vp, sz, vp]),
        "rs_extract_postfix_batch": (i32, [vp, vp, sz, C.c_uint32, vp, vp, vp]),
        "rs_extract_postfix_batch_dev": (i32, [vp, vp, sz, C.c_uint32, vp, vp, vp, vp]),
        "rs_locate_batch": (i32, [vp, vp, i32, sz, u64, C.c_uint32, vp, u64, vp, vp, vp, vp]),
        "rs_locate_batch_dev": (i32, [vp, vp, i32, sz, u64, C.c_uint32, vp, u64, vp, vp, vp, vp, C.POINTER(u64), vp]),
        "rs_index_set_samples": (i32, [vp, vp, vp, u64, i32]),
        "rs_samples_batch": (i32, [vp, vp, sz, vp, u64, vp]),
        "rs_carriers_batch": (i32, [vp, vp, vp, sz, vp, u64, vp]),
        "rs_find_interval_batch": (i32, [vp, vp, i32, sz, vp, vp]),
        "rs_find_interval_var_batch": (i32, [vp, vp, vp, sz, vp, vp]),
        "rs_find_interval_batch_dev": (i32, [vp, vp, i32, sz, vp, vp, vp]),
        "rs_find_interval_var_batch_dev": (i32, [vp, vp, vp, sz, vp, vp, vp]),
        "rs_count_batch": (i32, [vp, vp, i32, sz, i32, vp]),
        "rs_count_batch_dev": (i32, [vp, vp, i32, sz, i32, vp, vp]),
        "rs_count_batch_noalloc_dev": (i32, [vp, vp, i32, sz, i32, vp, vp]),
        "rs_index_set_sai": (i32, [vp, vp, u64]),
        "rs_extract_prefix_batch": (i32, [vp, vp, sz, C.c_uint32, vp, vp, vp, vp]),
        "rs_extract_prefix_batch_dev": (i32, [vp, vp, sz, C.c_uint32, vp, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _LIB = L
    return L


NOALLOC_MAX_QUERIES = 1 << 20  # RS_NOALLOC_MAX_QUERIES, include/readserver_b200.h


def unpack_hits40(raw: np.ndarray, n: int) -> np.ndarray:
    """The first n hits of an rs_pop_query_hits40 list (5 bytes each) in rs_pop_query's 64-bit form."""
    b = np.ascontiguousarray(raw[: 5 * n], dtype=np.uint8).reshape(n, 5).astype(np.uint64)
    v = b[:, 0] | (b[:, 1] << np.uint64(8)) | (b[:, 2] << np.uint64(16)) | (b[:, 3] << np.uint64(24)) | (b[:, 4] << np.uint64(32))
    out = ((v >> np.uint64(32)) << np.uint64(40)) | (v & np.uint64(0xFFFFFFFF))
    out[v == np.uint64((1 << 40) - 1)] = np.iinfo(np.uint64).max
    return out


def _check(rc: int):
    if rc != 0:
        raise RsError(rc, (lib().rs_last_error() or b"").decode("utf-8", "replace"))


def measure_random_access(device: int = 0, buffer_bytes: int = 16 << 30, lines: int = 1 << 28, granule: int = 128) -> float:
    """GB/s of random aligned loads of `granule` bytes over a buffer much larger than L2 (roofline denominator)."""
    v = C.c_double()
    _check(lib().rs_measure_random_access_granule(device, buffer_bytes, lines, granule, C.byref(v)))
    return float(v.value)


def _ptr(a) -> int:
    """Address of a numpy array / torch tensor / None."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _ascii(a) -> np.ndarray:
    if isinstance(a, str):
        a = a.encode()
    if isinstance(a, (bytes, bytearray)):
        return np.frombuffer(bytes(a), dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


class GpuBWT:
    """One BWT shard resident in the HBM of one B200 (``RLEBWT``, rlebwt.cpp:11-32)."""

    def __init__(self, bwt_path: str | None = None, *, runs=None, num_strings: int = 0, num_symbols: int = 0,
                 device: int = 0, encoded_path: str | None = None, host_encoder: bool = False):
        """bwt_path: an SGA .bwt file; runs: its run bytes (numpy: host memory; torch CUDA tensor: device memory,
        encoded in place on the GPU); encoded_path: a file written by save(); host_encoder: build the blocks with
        the host-side encoder instead of the device one (cross-check)."""
        L = lib()
        h = C.c_void_p()
        if encoded_path is not None:
            _check(L.rs_index_open_encoded(os.fsencode(encoded_path), device, C.byref(h)))
        elif bwt_path is not None:
            _check(L.rs_index_open(os.fsencode(bwt_path), device, C.byref(h)))
        elif not isinstance(runs, np.ndarray) and hasattr(runs, "data_ptr"):
            assert runs.is_cuda and runs.is_contiguous() and runs.element_size() == 1
            _check(L.rs_index_from_runs_dev(runs.data_ptr(), runs.numel(), num_strings, num_symbols, device, C.byref(h)))
        else:
            runs = np.ascontiguousarray(runs, dtype=np.uint8)
            f = L.rs_index_from_runs_host if host_encoder else L.rs_index_from_runs
            _check(f(runs.ctypes.data, runs.size, num_strings, num_symbols, device, C.byref(h)))
        self._h = h
        self.device = device
        info = Info()
        _check(L.rs_index_get_info(h, C.byref(info)))
        self.info = info

    # -- lifecycle -------------------------------------------------------------------------------
    def close(self):
        h = getattr(self, "_h", None)
        if h:
            lib().rs_index_close(h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_tail_emulation(self, on: bool):
        _check(lib().rs_index_set_tail_emulation(self._h, int(on)))

    def set_prefix_table(self, k: int):
        """(Re)build the k-mer suffix table with k characters; 0 drops it. Results do not depend on it."""
        _check(lib().rs_index_set_prefix_table(self._h, k))
        _check(lib().rs_index_get_info(self._h, C.byref(self.info)))

    def set_dense(self, on: bool):
        """Build / drop the dense rank blocks (0.5 bytes per symbol; one memory access per rank). Results do not change."""
        _check(lib().rs_index_set_dense(self._h, int(on)))
        _check(lib().rs_index_get_info(self._h, C.byref(self.info)))

    def enable_stats(self, on: bool = True):
        _check(lib().rs_index_enable_stats(self._h, int(on)))

    def read_stats(self, reset: bool = True) -> tuple[int, int]:
        """(updateInterval steps, LF-walk steps) executed by the *_dev calls since the last reset."""
        a, b = C.c_uint64(), C.c_uint64()
        _check(lib().rs_index_read_stats(self._h, C.byref(a), C.byref(b), int(reset)))
        return int(a.value), int(b.value)

    def set_sai(self, sai):
        if not isinstance(sai, np.ndarray) and hasattr(sai, "data_ptr"):  # torch CUDA tensor (int64 / uint64)
            assert sai.is_cuda and sai.is_contiguous() and sai.element_size() == 8
            _check(lib().rs_index_set_sai_dev(self._h, sai.data_ptr(), sai.numel()))
            return
        sai = np.ascontiguousarray(sai, dtype=np.uint64)
        _check(lib().rs_index_set_sai(self._h, sai.ctypes.data, sai.size))

    def save(self, path: str):
        """Persist the HBM-layout index (the .bpi2 analogue): reopen with GpuBWT(encoded_path=...)."""
        _check(lib().rs_index_save(self._h, os.fsencode(path)))

    def export(self, what: int) -> np.ndarray:
        """Raw bytes of one device array: 0 blocks, 1 position directory, 2..6 select directories, 7 suffix table."""
        n = C.c_uint64()
        _check(lib().rs_index_export(self._h, what, None, 0, C.byref(n)))
        buf = np.zeros(max(int(n.value), 1), dtype=np.uint8)
        _check(lib().rs_index_export(self._h, what, buf.ctypes.data, buf.size, C.byref(n)))
        return buf[: int(n.value)]

    # -- class BWT -------------------------------------------------------------------------------
    def get_bwlen(self) -> int:
        return int(lib().rs_get_bwlen(self._h))

    def get_pc(self, b) -> int:
        b = b.encode() if isinstance(b, str) else bytes([b]) if isinstance(b, int) else b
        return int(lib().rs_get_pc(self._h, b))

    def get_occ(self, b, index) -> np.ndarray:
        b = _ascii(b)
        index = np.ascontiguousarray(index, dtype=np.uint64)
        if b.size == 1 and index.size != 1:
            b = np.full(index.size, b[0], dtype=np.uint8)
        assert b.size == index.size
        out = np.empty(index.size, dtype=np.uint64)
        _check(lib().rs_get_occ_batch(self._h, b.ctypes.data, index.ctypes.data, index.size, out.ctypes.data))
        return out

    def get_char(self, index) -> np.ndarray:
        index = np.ascontiguousarray(index, dtype=np.uint64)
        out = np.empty(index.size, dtype=np.uint8)
        _check(lib().rs_get_char_batch(self._h, index.ctypes.data, index.size, out.ctypes.data))
        return out

    def get_f(self, index) -> np.ndarray:
        index = np.ascontiguousarray(index, dtype=np.uint64)
        out = np.empty(index.size, dtype=np.uint8)
        _check(lib().rs_get_f_batch(self._h, index.ctypes.data, index.size, out.ctypes.data))
        return out

    def get_occ_at(self, b, count) -> np.ndarray:
        b = _ascii(b)
        count = np.ascontiguousarray(count, dtype=np.uint64)
        if b.size == 1 and count.size != 1:
            b = np.full(count.size, b[0], dtype=np.uint8)
        assert b.size == count.size
        out = np.empty(count.size, dtype=np.uint64)
        _check(lib().rs_get_occ_at_batch(self._h, b.ctypes.data, count.ctypes.data, count.size, out.ctypes.data))
        return out

    # -- query.h ---------------------------------------------------------------------------------
    def find_interval(self, kmers, k: int | None = None):
        """findInterval for a [Q, k] uint8 ASCII array (or a single str). Returns (lower, upper) uint64."""
        if isinstance(kmers, str):
            k = len(kmers)
        kmers = _ascii(kmers)
        if k is None:
            k = kmers.shape[-1]
        flat = kmers.reshape(-1)
        if k < 1 or flat.size % k:
            raise RsError(5, "k-mer array size is not a multiple of k")
        q = flat.size // k
        lo = np.empty(q, dtype=np.uint64)
        up = np.empty(q, dtype=np.uint64)
        _check(lib().rs_find_interval_batch(self._h, flat.ctypes.data, k, q, lo.ctypes.data, up.ctypes.data))
        return lo, up

    def find_interval_var(self, strings: list[str]):
        text = np.frombuffer("".join(strings).encode(), dtype=np.uint8)
        offs = np.zeros(len(strings) + 1, dtype=np.uint64)
        np.cumsum([len(s) for s in strings], out=offs[1:])
        lo = np.empty(len(strings), dtype=np.uint64)
        up = np.empty(len(strings), dtype=np.uint64)
        text = np.ascontiguousarray(text) if text.size else np.zeros(1, dtype=np.uint8)
        _check(lib().rs_find_interval_var_batch(self._h, text.ctypes.data, offs.ctypes.data, len(strings),
                                                lo.ctypes.data, up.ctypes.data))
        return lo, up

    def count(self, kmers, k: int | None = None, strand: int = 0) -> np.ndarray:
        """count_reads (service.cpp:279-315): matching rows per k-mer on one strand."""
        kmers = _ascii(kmers)
        if k is None:
            k = kmers.shape[-1]
        flat = kmers.reshape(-1)
        q = flat.size // k
        out = np.empty(q, dtype=np.uint64)
        _check(lib().rs_count_batch(self._h, flat.ctypes.data, k, q, strand, out.ctypes.data))
        return out

    def extract_prefix(self, rows, max_steps: int = 1024, want_bases: bool = False, want_read_index: bool = False):
        """extractPrefix (query.cpp:43-63) per row: (lexrank, steps[, read_index][, bases [n, max_steps]])."""
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        n = rows.size
        lex = np.empty(n, dtype=np.uint64)
        st = np.empty(n, dtype=np.uint32)
        ri = np.empty(n, dtype=np.uint64) if want_read_index else None
        bases = np.zeros((n, max_steps), dtype=np.uint8) if want_bases else None
        _check(lib().rs_extract_prefix_batch(self._h, rows.ctypes.data, n, max_steps, lex.ctypes.data, st.ctypes.data,
                                             _ptr(ri), _ptr(bases)))
        out = [lex, st]
        if want_read_index:
            out.append(ri)
        if want_bases:
            out.append(bases)
        return tuple(out)

    def extract_postfix(self, rows, max_steps: int = 1024, want_bases: bool = False):
        """extractPostfix (query.cpp:65-85) per row: (read_index, steps[, bases [n, max_steps]])."""
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        n = rows.size
        ri = np.empty(n, dtype=np.uint64)
        st = np.empty(n, dtype=np.uint32)
        bases = np.zeros((n, max_steps), dtype=np.uint8) if want_bases else None
        _check(lib().rs_extract_postfix_batch(self._h, rows.ctypes.data, n, max_steps, ri.ctypes.data, st.ctypes.data,
                                              _ptr(bases)))
        return (ri, st, bases) if want_bases else (ri, st)

    def extract_read(self, rows, max_steps: int = 1024) -> list[str]:
        """extractPrefix(row) + extractPostfix(row): the read each row belongs to (query(), query.cpp:93-97)."""
        _, ps, pb = self.extract_prefix(rows, max_steps, want_bases=True)
        _, qs, qb = self.extract_postfix(rows, max_steps, want_bases=True)
        out = []
        for i in range(len(ps)):
            if ps[i] == 0xFFFFFFFF or qs[i] == 0xFFFFFFFF:
                out.append(None)
            else:
                out.append(bytes(pb[i, :ps[i]]).decode() + bytes(qb[i, :qs[i]]).decode())
        return out

    # -- count + locate, sample sets -------------------------------------------------------------
    def locate(self, kmers, k: int | None = None, max_hits: int = 100_000, max_steps: int = 1024,
               want_read_index: bool = True):
        """Per k-mer: the rows of its interval (<= max_hits), each walked to its read (query(), query.cpp:87-100,
        without rebuilding strings).  Returns dict(hit_offsets, rows, lexrank, steps[, read_index])."""
        kmers = _ascii(kmers)
        if k is None:
            k = kmers.shape[-1]
        flat = kmers.reshape(-1)
        q = flat.size // k
        off = np.empty(q + 1, dtype=np.uint64)
        L = lib()
        rc = L.rs_locate_batch(self._h, flat.ctypes.data, k, q, max_hits, max_steps, off.ctypes.data, 0, None, None, None, None)
        if rc not in (0, 9):
            _check(rc)
        h = int(off[q])
        rows = np.empty(h, dtype=np.uint64)
        lex = np.empty(h, dtype=np.uint64)
        st = np.empty(h, dtype=np.uint32)
        ri = np.empty(h, dtype=np.uint64) if want_read_index else None
        if h:
            _check(L.rs_locate_batch(self._h, flat.ctypes.data, k, q, max_hits, max_steps, off.ctypes.data, h,
                                     rows.ctypes.data, lex.ctypes.data, _ptr(ri), st.ctypes.data))
        out = dict(hit_offsets=off, rows=rows, lexrank=lex, steps=st)
        if want_read_index:
            out["read_index"] = ri
        return out

    def set_samples(self, offsets, codes, size_of_sample: int, meta=None):
        """Sample table: read r owns codes [offsets[r], offsets[r+1]); meta (optional) = 2 bytes per code, the count
        and lane characters of a store with has_other_meta_data.  torch CUDA tensors are copied device to device."""
        self._sos = size_of_sample
        if not isinstance(offsets, np.ndarray) and hasattr(offsets, "data_ptr"):
            assert meta is None and offsets.is_cuda and offsets.is_contiguous() and codes.is_contiguous()
            _check(lib().rs_index_set_samples_dev(self._h, offsets.data_ptr(), codes.data_ptr(), offsets.numel() - 1,
                                                  size_of_sample))
            return
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        codes = np.ascontiguousarray(codes, dtype=np.uint8).reshape(-1)
        if meta is not None:
            meta = np.ascontiguousarray(meta, dtype=np.uint8).reshape(-1)
            assert meta.size == 2 * (codes.size // size_of_sample)
        _check(lib().rs_index_set_samples_meta(self._h, offsets.ctypes.data, codes.ctypes.data if codes.size else None,
                                               meta.ctypes.data if meta is not None and meta.size else None,
                                               offsets.size - 1, size_of_sample))

    def samples_meta(self, read_index):
        """(offsets [n+1], codes [total, sos], meta [total, 2]) for n reads of a table with count / lane metadata."""
        ri = np.ascontiguousarray(read_index, dtype=np.uint64)
        off = np.empty(ri.size + 1, dtype=np.uint64)
        L = lib()
        rc = L.rs_samples_meta_batch(self._h, ri.ctypes.data, ri.size, off.ctypes.data, 0, None, None)
        if rc not in (0, 9):
            _check(rc)
        tot = int(off[ri.size])
        codes = np.empty((tot, self._sos), dtype=np.uint8)
        meta = np.empty((tot, 2), dtype=np.uint8)
        if tot:
            _check(L.rs_samples_meta_batch(self._h, ri.ctypes.data, ri.size, off.ctypes.data, tot, codes.ctypes.data,
                                           meta.ctypes.data))
        return off, codes, meta

    def load_samples_file(self, path: str, size_of_sample: int, has_other_meta_data: bool = False):
        """Sample table from the ETL's merged `read / codes` file (load_data_into_rocksdb.cpp:41-49)."""
        _check(lib().rs_index_load_samples_file(self._h, path.encode(), size_of_sample, int(has_other_meta_data)))
        self._sos = size_of_sample

    def set_reads(self, reads: list):
        """Attach the packed read table: the reads in builder input order (ACGT only)."""
        bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
        text = np.frombuffer(b"".join(bs) or b"\0", dtype=np.uint8)
        off = np.zeros(len(bs) + 1, dtype=np.uint64)
        off[1:] = np.cumsum([len(b) for b in bs])
        _check(lib().rs_index_set_reads(self._h, text.ctypes.data, off.ctypes.data, len(bs)))

    def set_reads_arrays(self, text: np.ndarray, offsets: np.ndarray):
        """Packed read table from the concatenated ASCII reads (uint8) and their offsets (uint64, n + 1)."""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        _check(lib().rs_index_set_reads(self._h, text.ctypes.data, offsets.ctypes.data, offsets.size - 1))

    def reads_dev(self, d_read_index, n: int, max_len: int, d_out, d_out_len, stream: int = 0):
        _check(lib().rs_reads_batch_dev(self._h, _ptr(d_read_index), n, max_len, _ptr(d_out), _ptr(d_out_len), stream))

    def load_reads_file(self, path: str):
        _check(lib().rs_index_load_reads_file(self._h, path.encode()))

    def reads(self, read_index, max_len: int = 128) -> list:
        """Read strings by read index (None for an index outside the table)."""
        ri = np.ascontiguousarray(read_index, dtype=np.uint64)
        out = np.zeros((ri.size, max_len), dtype=np.uint8)
        ln = np.zeros(ri.size, dtype=np.uint32)
        _check(lib().rs_reads_batch(self._h, ri.ctypes.data, ri.size, max_len, out.ctypes.data, ln.ctypes.data))
        return [None if l == 0xFFFFFFFF else out[i, :min(int(l), max_len)].tobytes().decode() for i, l in enumerate(ln)]

    def samples(self, read_index):
        """(offsets [n+1], codes [total, size_of_sample]) for n reads, order and duplicates preserved."""
        ri = np.ascontiguousarray(read_index, dtype=np.uint64)
        off = np.empty(ri.size + 1, dtype=np.uint64)
        L = lib()
        rc = L.rs_samples_batch(self._h, ri.ctypes.data, ri.size, off.ctypes.data, 0, None)
        if rc not in (0, 9):
            _check(rc)
        tot = int(off[ri.size])
        codes = np.empty((tot, self._sos), dtype=np.uint8)
        if tot:
            _check(L.rs_samples_batch(self._h, ri.ctypes.data, ri.size, off.ctypes.data, tot, codes.ctypes.data))
        return off, codes

    def carriers(self, hit_offsets, read_index):
        """Per k-mer the sorted distinct sample codes over its hit reads: (offsets [Q+1], codes [total, sos])."""
        ho = np.ascontiguousarray(hit_offsets, dtype=np.uint64)
        ri = np.ascontiguousarray(read_index, dtype=np.uint64)
        q = ho.size - 1
        off = np.empty(q + 1, dtype=np.uint64)
        L = lib()
        rc = L.rs_carriers_batch(self._h, ho.ctypes.data, ri.ctypes.data if ri.size else None, q, off.ctypes.data, 0, None)
        if rc not in (0, 9):
            _check(rc)
        tot = int(off[q])
        codes = np.empty((tot, self._sos), dtype=np.uint8)
        if tot:
            _check(L.rs_carriers_batch(self._h, ho.ctypes.data, ri.ctypes.data, q, off.ctypes.data, tot, codes.ctypes.data))
        return off, codes

    # -- device-pointer forms (torch tensors on self.device; asynchronous on `stream`) -------------
    def find_interval_dev(self, d_kmers, k: int, q: int, d_lower, d_upper, stream: int = 0):
        _check(lib().rs_find_interval_batch_dev(self._h, _ptr(d_kmers), k, q, _ptr(d_lower), _ptr(d_upper), stream))

    def tile_intervals(self, queries: list, kmer: int, skip: int = 0):
        """get_tiles + findInterval per tile (service.cpp:232-246, 466-500) for a batch of query strings.
        Returns (tile_offsets [Q+1], lower [T], upper [T]); tiles the reference skips (non-ACGT) are (1, 0)."""
        bs = [q.encode() if isinstance(q, str) else bytes(q) for q in queries]
        text = np.frombuffer(b"".join(bs) or b"\0", dtype=np.uint8)
        offs = np.zeros(len(bs) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(b) for b in bs])
        step = skip + 1
        cap = int(sum(0 if len(b) < kmer else (len(b) - kmer) // step + 1 for b in bs))
        toff = np.zeros(len(bs) + 1, dtype=np.uint64)
        lo = np.empty(max(cap, 1), dtype=np.uint64)
        up = np.empty(max(cap, 1), dtype=np.uint64)
        _check(lib().rs_tile_intervals_batch(self._h, text.ctypes.data, offs.ctypes.data, len(bs), kmer, skip,
                                             toff.ctypes.data, cap, lo.ctypes.data, up.ctypes.data))
        return toff, lo[:cap], up[:cap]

    def build_locate_table(self, max_read_len: int = 4096, every: int = 1):
        """every = 1: 8 bytes per BWT symbol of HBM, the locate stage reads one entry per hit row instead of walking;
        every > 1: sampled (n (1/8 + 1/64 + 8/every) bytes), a hit row is walked fewer than `every` steps."""
        _check(lib().rs_index_build_locate_samples(self._h, max_read_len, every))
        _check(lib().rs_index_get_info(self._h, C.byref(self.info)))

    def drop_locate_table(self):
        _check(lib().rs_index_drop_locate_table(self._h))
        _check(lib().rs_index_get_info(self._h, C.byref(self.info)))

    def gt_reads(self, windows: list, pos, kmer: int, skip: int = 0, max_interval_size: int = 10000,
                 max_steps: int = 1024, max_rows: int = 1 << 26):
        """Seed stage of /gt, find_gt_reads over every tile (service.cpp:507-711, 1044-1072), for a batch of
        oriented windows with their 1-based site positions.  Returns (offsets [Q+1], read_index, flags [Q],
        candidate_rows): each window's sorted unique read indices and its GT_FLAG_* bits."""
        bs = [w.encode() if isinstance(w, str) else bytes(w) for w in windows]
        text = np.frombuffer(b"".join(bs) or b"\0", dtype=np.uint8)
        offs = np.zeros(len(bs) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(b) for b in bs])
        pos = np.ascontiguousarray(pos, dtype=np.uint64)
        assert pos.size == len(bs)
        out_off = np.zeros(len(bs) + 1, dtype=np.uint64)
        flags = np.zeros(max(len(bs), 1), dtype=np.uint32)
        cand = C.c_uint64()
        cap = 1 << 16
        while True:
            ri = np.empty(cap, dtype=np.uint64)
            rc = lib().rs_gt_reads_batch(self._h, text.ctypes.data, offs.ctypes.data, pos.ctypes.data, len(bs), kmer, skip,
                                         max_interval_size, max_steps, max_rows, out_off.ctypes.data, cap, ri.ctypes.data,
                                         flags.ctypes.data, C.byref(cand))
            if rc == 9 and int(cand.value) <= max_rows and int(out_off[-1]) > cap:  # RS_STATUS_CAPACITY on the output
                cap = int(out_off[-1])
                continue
            _check(rc)
            break
        return out_off, ri[:int(out_off[-1])], flags[:len(bs)], int(cand.value)

    def count_dev(self, d_kmers, k: int, q: int, strand: int, d_counts, stream: int = 0):
        _check(lib().rs_count_batch_dev(self._h, _ptr(d_kmers), k, q, strand, _ptr(d_counts), stream))

    def count_noalloc_dev(self, d_kmers, k: int, q: int, strand: int, d_counts, stream: int = 0):
        """count_dev for callers that may not allocate: scratch in the library's static device array
        (q <= NOALLOC_MAX_QUERIES, k <= 31; calls overlapping on the GPU race -- include/readserver_b200.h)."""
        _check(lib().rs_count_batch_noalloc_dev(self._h, _ptr(d_kmers), k, q, strand, _ptr(d_counts), stream))

    def locate_dev(self, d_kmers, k: int, q: int, max_hits: int, max_steps: int, d_hit_offsets, capacity: int,
                   d_rows=None, d_lexrank=None, d_read_index=None, d_steps=None, stream: int = 0) -> int:
        """count + locate on device buffers; returns the total number of hits (raises RsError(9) if > capacity)."""
        tot = C.c_uint64()
        _check(lib().rs_locate_batch_dev(self._h, _ptr(d_kmers), k, q, max_hits, max_steps, _ptr(d_hit_offsets), capacity,
                                         _ptr(d_rows), _ptr(d_lexrank), _ptr(d_read_index), _ptr(d_steps), C.byref(tot),
                                         stream))
        return int(tot.value)

    def extract_prefix_dev(self, d_rows, n: int, max_steps: int, d_lexrank=None, d_steps=None, d_read_index=None,
                           d_bases=None, stream: int = 0):
        _check(lib().rs_extract_prefix_batch_dev(self._h, _ptr(d_rows), n, max_steps, _ptr(d_lexrank), _ptr(d_steps),
                                                 _ptr(d_read_index), _ptr(d_bases), stream))


class ShardSet:
    """The BWT shards resident on one GPU, queried together (rs_shardset_* / rs_pop_query of the C ABI): what the
    reference's front end does over its per-shard workers -- broadcast, sum of counts, concatenation of read lists
    (src/service/server.cpp:124, 184-197, 404-410) -- as one device pipeline."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        self.device = device
        self._keep = []
        _check(lib().rs_shardset_create(device, C.byref(self._h)))

    def add(self, bwt: "GpuBWT", shard_id: int):
        _check(lib().rs_shardset_add(self._h, bwt._h, shard_id))
        self._keep.append(bwt)  # the set borrows the index

    def seal(self, max_read_len: int, locate_every: int = 1):
        """locate_every = 1: read-index tables (4 bytes per BWT row); > 1: sampled locate tables; 0: plain walks."""
        _check(lib().rs_shardset_seal(self._h, max_read_len, locate_every))
        return self.info

    @property
    def info(self) -> ShardSetInfo:
        i = ShardSetInfo()
        _check(lib().rs_shardset_get_info(self._h, C.byref(i)))
        return i

    def enable_stats(self, on: bool = True):
        _check(lib().rs_shardset_enable_stats(self._h, int(on)))

    def read_stats(self, reset: bool = True) -> tuple[int, int, int]:
        """(updateInterval steps, LF-walk steps, 64-byte blocks read by K1) since the last reset."""
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib().rs_shardset_read_stats(self._h, C.byref(a), C.byref(b), C.byref(c), int(reset)))
        return int(a.value), int(b.value), int(c.value)

    def enable_timing(self, on: bool = True):
        _check(lib().rs_shardset_enable_timing(self._h, int(on)))

    def read_timing(self):
        """({stage: ms}, {stage: kernel launches}) since the last read; stages pack, search (K1), offsets, locate."""
        ms = (C.c_double * 4)()
        n = (C.c_uint64 * 4)()
        _check(lib().rs_shardset_read_timing(self._h, ms, n))
        names = ("pack", "search", "offsets", "locate")
        return {k: float(v) for k, v in zip(names, ms)}, {k: int(v) for k, v in zip(names, n)}

    def query_dev(self, d_kmers, k: int, q: int, packed: bool, max_hits: int, d_counts, d_hit_offsets=None, d_hits=None,
                  capacity: int = 0, d_carriers=None, d_total=None, stream: int = 0, strand: int = 0):
        """Asynchronous on `stream`, device buffers (torch tensors); see rs_pop_query_dev."""
        _check(lib().rs_pop_query_dev(self._h, _ptr(d_kmers), k, q, int(packed), strand, max_hits, _ptr(d_counts), _ptr(d_hit_offsets),
                                      _ptr(d_hits), capacity, _ptr(d_carriers), _ptr(d_total), stream))

    def query(self, kmers, k: int | None = None, max_hits: int = 1 << 20, locate: bool = True, carriers: bool = False,
              packed: bool = False, strand: int = 0, hits40: bool = False):
        """Host arrays in, host arrays out: dict(counts, hit_offsets, hits, carriers).  hits40: the list travels as 5
        bytes per hit (rs_pop_query_hits40) and is widened here."""
        if packed:
            kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
            q = kmers.size
        else:
            kmers = _ascii(kmers)
            if k is None:
                k = kmers.shape[-1]
            kmers = kmers.reshape(-1)
            q = kmers.size // k
        counts = np.zeros(q, dtype=np.uint64)
        out = {"counts": counts}
        L = lib()
        tot = C.c_uint64()
        if not locate:
            _check(L.rs_pop_query(self._h, kmers.ctypes.data, k, q, int(packed), strand, max_hits, counts.ctypes.data, None, None, 0,
                                  None, C.byref(tot)))
            return out
        off = np.zeros(q + 1, dtype=np.uint64)
        rc = L.rs_pop_query(self._h, kmers.ctypes.data, k, q, int(packed), strand, max_hits, counts.ctypes.data, off.ctypes.data, None,
                            0, None, C.byref(tot))
        if rc not in (0, 9):
            _check(rc)
        cap = int(tot.value)
        W = int(self.info.carrier_words)
        carr = np.zeros((q, W), dtype=np.uint32) if carriers else None
        if hits40:
            raw = np.zeros(5 * max(cap, 1), dtype=np.uint8)
            _check(L.rs_pop_query_hits40(self._h, kmers.ctypes.data, k, q, int(packed), strand, max_hits, counts.ctypes.data,
                                         off.ctypes.data, raw.ctypes.data, cap, _ptr(carr), C.byref(tot)))
            hits = unpack_hits40(raw, cap)
        else:
            hits = np.zeros(max(cap, 1), dtype=np.uint64)
            _check(L.rs_pop_query(self._h, kmers.ctypes.data, k, q, int(packed), strand, max_hits, counts.ctypes.data, off.ctypes.data,
                                  hits.ctypes.data, cap, _ptr(carr), C.byref(tot)))
        out.update(hit_offsets=off, hits=hits[:cap])
        if carriers:
            out["carriers"] = carr
        return out

    def close(self):
        if getattr(self, "_h", None):
            lib().rs_shardset_destroy(self._h)
            self._h = None
        self._keep = []

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class PopComm:
    """Result buffers of one rank for population queries across the GPUs of a box (rs_popcomm_* of the C ABI): counts,
    hit lists and carrier sets are fanned in to the k-mer's owner through NVLink peer memory by the pipeline's own
    kernels (the front end's sum / concatenation over shard replies, src/service/server.cpp:184-197, 404-410)."""
    HANDLE_BYTES = 64

    def __init__(self, device: int, rank: int, world: int, max_queries: int, hit_capacity: int):
        self._h = C.c_void_p()
        self.device, self.rank, self.world = device, rank, world
        self.max_queries, self.hit_capacity = max_queries, hit_capacity
        _check(lib().rs_popcomm_create(device, rank, world, max_queries, hit_capacity, C.byref(self._h)))

    def export(self) -> np.ndarray:
        h = np.zeros(self.HANDLE_BYTES, dtype=np.uint8)
        _check(lib().rs_popcomm_export(self._h, h.ctypes.data))
        return h

    def connect(self, handles: np.ndarray):
        handles = np.ascontiguousarray(handles, dtype=np.uint8).reshape(-1)
        assert handles.size == self.world * self.HANDLE_BYTES
        _check(lib().rs_popcomm_connect(self._h, handles.ctypes.data))

    def query_dev(self, shardset: "ShardSet", d_kmers, k: int, q: int, packed: bool, max_hits: int, want_hits: bool = True,
                  want_carriers: bool = False, stream: int = 0, strand: int = 0):
        _check(lib().rs_pop_query_comm_dev(shardset._h, self._h, _ptr(d_kmers), k, q, int(packed), strand, max_hits, int(want_hits),
                                           int(want_carriers), stream))

    def results(self, q: int, device=None):
        """torch views (no copy) of this rank's result buffers for a batch of q k-mers:
        dict(counts [q], hit_offsets [q+1], hits [capacity], carriers [q, 4] int32, total [1])."""
        import torch
        p = [C.c_void_p() for _ in range(5)]
        cap = C.c_uint64()
        _check(lib().rs_popcomm_results(self._h, *[C.byref(x) for x in p], C.byref(cap)))
        dev = device if device is not None else f"cuda:{self.device}"

        def view(ptr, shape, typestr):
            class _A:
                __cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr.value), False), "version": 3}
            return torch.as_tensor(_A(), device=dev)
        return dict(counts=view(p[0], (q,), "<i8"), hit_offsets=view(p[1], (q + 1,), "<i8"),
                    hits=view(p[2], (max(int(cap.value), 1),), "<i8"), carriers=view(p[3], (q, 4), "<i4"),
                    total=view(p[4], (1,), "<i8"))

    def timed_out(self) -> bool:
        v = C.c_int32()
        _check(lib().rs_popcomm_error(self._h, C.byref(v)))
        return bool(v.value)

    def close(self):
        if getattr(self, "_h", None):
            lib().rs_popcomm_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def pack_kmers(kmers: np.ndarray) -> np.ndarray:
    """[Q, k] ASCII -> uint64 keys of the packed-query entry points: 2 bits per base (A,C,G,T = 0..3), first base
    most significant; bit 63 set when the k-mer holds any other byte (it then counts 0, service.cpp:299-301)."""
    kmers = np.ascontiguousarray(kmers, dtype=np.uint8)
    lut = np.full(256, 255, dtype=np.uint8)
    for i, c in enumerate(b"ACGT"):
        lut[c] = i
    codes = lut[kmers]
    bad = (codes == 255).any(axis=1)
    v = np.zeros(kmers.shape[0], dtype=np.uint64)
    for j in range(kmers.shape[1]):
        v = (v << np.uint64(2)) | (codes[:, j] & 3).astype(np.uint64)
    v[bad] |= np.uint64(1) << np.uint64(63)
    return v


class PeerComm:
    """Fan-in of per-shard counts across the GPUs of one box through NVLink peer memory (rs_comm_* of the
    C ABI; replaces the front end's sum over shard replies, src/service/server.cpp:184-197, 404-410).
    The caller exchanges the 64-byte handles (e.g. with torch.distributed.all_gather)."""
    HANDLE_BYTES = 64

    def __init__(self, device: int, rank: int, world: int, max_queries: int):
        self._h = C.c_void_p()
        self.rank, self.world, self.max_queries = rank, world, max_queries
        _check(lib().rs_comm_create(device, rank, world, max_queries, C.byref(self._h)))

    def export(self) -> np.ndarray:
        h = np.zeros(self.HANDLE_BYTES, dtype=np.uint8)
        _check(lib().rs_comm_export(self._h, h.ctypes.data))
        return h

    def connect(self, handles: np.ndarray):
        handles = np.ascontiguousarray(handles, dtype=np.uint8).reshape(-1)
        assert handles.size == self.world * self.HANDLE_BYTES
        _check(lib().rs_comm_connect(self._h, handles.ctypes.data))

    def begin(self, q: int, stream: int = 0):
        _check(lib().rs_comm_begin(self._h, q, stream))

    def accumulate(self, bwt: "GpuBWT", d_kmers, k: int, q: int, strand: int = 0, stream: int = 0):
        _check(lib().rs_count_accumulate_dev(bwt._h, self._h, _ptr(d_kmers), k, q, strand, stream))

    def finish(self, stream: int = 0) -> int:
        """Barrier on the stream; returns the device address of this step's Q u64 counts."""
        p = C.c_void_p()
        _check(lib().rs_comm_finish(self._h, stream, C.byref(p)))
        return int(p.value)

    def timed_out(self) -> bool:
        v = C.c_int32()
        _check(lib().rs_comm_error(self._h, C.byref(v)))
        return bool(v.value)

    def close(self):
        if self._h:
            lib().rs_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
