"""TEST INFRASTRUCTURE ONLY: CPU restatement of the seed stage of /gt.

``find_gt_reads`` follows the reference's function of that name statement by statement
(/root/reference/src/service/service.cpp:507-711) on top of the oracle's findInterval / extractPrefix /
extractPostfix (the reference's own, through ``oracle/_ref``, when built).  ``gt_task_reads`` is the tile loop of
``GtTask::run`` (service.cpp:1044-1072).  service.cpp itself does not compile here (ZeroMQ, protobuf, RocksDB),
so this restatement is anchored on the reference's primitives, not on a run of the function itself: parity for
this row is "unpinned" beyond them.

All arithmetic is size_t (64-bit wrapping), as in the reference.  Where the reference leaves defined behaviour
``GtUndefined`` is raised: ``throw`` when ``w.substr(start-1, ..)`` is called with start = 0
(std::out_of_range), ``hang`` when an extension loop searches the same (truncated) string twice -- the interval
cannot shrink any more and the loop never ends.
"""
from __future__ import annotations

import numpy as np

M64 = (1 << 64) - 1
INDEL_ALLOWANCE = 4          # service.cpp:509
MAX_INTERVAL_SIZE = 10000    # service.cpp:91


class GtUndefined(Exception):
    def __init__(self, kind: str):
        super().__init__(kind)
        self.kind = kind


def _substr(w: str, a: int, n: int) -> str:
    """std::string::substr: throws when a > size(), truncates at the end"""
    if a < 0 or a > len(w):
        raise GtUndefined("throw")
    return w[a:a + n]


def _wide(iv, max_iv: int) -> bool:
    return ((iv[1] - iv[0] + 1) & M64) > max_iv


class _Searcher:
    """findInterval with the hang check: a wide interval of a string equal to the previous one never changes"""

    def __init__(self, ix, max_iv):
        self.ix, self.max_iv, self.prev = ix, max_iv, None

    def first(self, kmer):
        self.prev = kmer
        return self.ix.find_interval(kmer)

    def again(self, kmer):
        if kmer == self.prev:
            raise GtUndefined("hang")
        self.prev = kmer
        return self.ix.find_interval(kmer)


def _rows(iv):
    return np.arange(iv[0], iv[1] + 1, dtype=np.uint64) if iv[1] >= iv[0] else np.zeros(0, dtype=np.uint64)


def find_gt_reads(ix, w: str, index: int, kmers: int, skip: int, pos: int, max_interval_size: int = MAX_INTERVAL_SIZE,
                  seeds: list | None = None) -> set:
    """Read indices (builder input order) of the reads find_gt_reads returns for tile `index` of window w.
    The reference returns the strings prefix+postfix; reads are unique in the index, so the set of their indices
    (terminal row of the postfix walk) is the same set.  `seeds`, when given, receives
    (lower, upper, kind, thr) of every interval whose rows were walked."""
    out: set = set()
    length = len(w)
    kmer = _substr(w, (skip + 1) * index, kmers)
    if any(ch not in "ACGT" for ch in kmer):          # :513-515
        return out
    S = _Searcher(ix, max_interval_size)
    interval = S.first(kmer)
    start = (skip + 1) * index + 1                    # 1-based
    end = start + kmers - 1

    def walk(iv, kind, thr):
        if seeds is not None:
            seeds.append((iv[0], iv[1], kind, thr))
        rows = _rows(iv)
        if not rows.size:
            return
        _, pre, ri, post = ix.locate_batch(rows)
        for j in range(rows.size):
            if kind == "postfix":                     # :533-537
                postfix_size = (int(post[j]) - kmers) & M64
                if thr > ((postfix_size + INDEL_ALLOWANCE) & M64):
                    continue
            elif kind == "prefix":                    # :606-609
                if thr > int(pre[j]) + INDEL_ALLOWANCE:
                    continue
            out.add(int(ri[j]))

    if pos > end:                                     # k-mer on the left, :522
        if _wide(interval, max_interval_size):
            count = 0
            while _wide(interval, max_interval_size):   # to the right, :527-532
                count += 1
                interval = S.again(_substr(w, start - 1, kmers + count))
                end += 1
            walk(interval, "postfix", (pos - end) & M64)
            if index != 0:                            # :546-577
                start = (skip + 1) * index + 1
                end = start + kmers - 1
                count = 1
                start -= 1
                interval = S.first(_substr(w, start - 1, kmers + count))
                while _wide(interval, max_interval_size):
                    count += 1
                    if start > 1:
                        start -= 1
                    else:
                        end += 1
                    interval = S.again(_substr(w, start - 1, kmers + count))
                walk(interval, "postfix", (pos - end) & M64)
        else:
            walk(interval, "postfix", (pos - end) & M64)
    elif pos < start:                                 # k-mer on the right, :591
        if _wide(interval, max_interval_size):
            count = 0
            while _wide(interval, max_interval_size):   # to the left, :598-603
                count += 1
                start -= 1
                interval = S.again(_substr(w, start - 1, kmers + count))
            walk(interval, "prefix", (start - pos) & M64)
            if end < length:                          # :616-646
                start = (skip + 1) * index + 1
                end = start + kmers - 1
                count = 1
                end += 1
                interval = S.first(_substr(w, start - 1, kmers + count))
                while _wide(interval, max_interval_size):
                    count += 1
                    if end < length:
                        end += 1
                    else:
                        start -= 1
                    interval = S.again(_substr(w, start - 1, kmers + count))
                walk(interval, "prefix", (start - pos) & M64)
        else:
            walk(interval, "prefix", (start - pos) & M64)
    else:                                             # k-mer over the site, :660
        if _wide(interval, max_interval_size):
            count = 0
            while _wide(interval, max_interval_size):
                count += 1
                interval = S.again(_substr(w, start - 1, kmers + count))
                end += 1
            walk(interval, "all", 0)
            start = (skip + 1) * index + 1
            end = start + kmers - 1
            count = 1
            start -= 1
            interval = S.first(_substr(w, start - 1, kmers + count))
            while _wide(interval, max_interval_size):
                count += 1
                start -= 1
                interval = S.again(_substr(w, start - 1, kmers + count))
            walk(interval, "all", 0)
        else:
            walk(interval, "all", 0)
    return out


def gt_task_reads(ix, w: str, pos: int, kmers: int, skip: int, max_interval_size: int = MAX_INTERVAL_SIZE):
    """The tile loop of GtTask::run (service.cpp:1044-1072) on an already oriented and cut window: returns
    (sorted read indices, undefined kinds met).  A tile on which the reference throws or hangs contributes
    nothing here and is reported instead."""
    if pos > len(w) or len(w) < kmers:
        return [], set()
    tiles = (len(w) - kmers) // (skip + 1) + 1
    reads: set = set()
    undefined: set = set()
    for i in range(tiles):
        try:
            reads |= find_gt_reads(ix, w, i, kmers, skip, pos, max_interval_size)
        except GtUndefined as e:
            undefined.add(e.kind)
    return sorted(reads), undefined
