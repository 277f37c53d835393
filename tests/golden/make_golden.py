"""Generates the golden vectors under tests/golden/ FROM THE UNMODIFIED REFERENCE.

Run in the build container (needs /root/reference):   python -m tests.golden.make_golden

The reference sources (src/bwt/{query,rlebwt,rlebwt_reader}.cpp) are compiled where they lie into
oracle/_ref/libref_bwt.so by oracle/Makefile; every expected value below is what that library
returned through its public `class BWT` / query.h interface.  The reference has no fixtures of its
own, so these vectors are what pins the oracle port and the CUDA path when
/root/reference is not there (the GPU box).

  small.bwt / small.npz      the 'small' synthetic read set (depth-2 tree, large tail region)
  clean.bwt / clean.npz      a read set whose tail region is empty (the reference is exact everywhere)
  deep_rand.npz              2^20+ random run bytes regenerated from a seed (depth-3 tree)
"""
from __future__ import annotations

import hashlib
import os
import tempfile

import numpy as np

import oracle
from tests import common

HERE = os.path.dirname(os.path.abspath(__file__))
SYMS = np.frombuffer(b"$ACGT", dtype=np.uint8)


def random_runs(seed: int, nruns: int) -> tuple[np.ndarray, int]:
    """Deterministic run bytes: adjacent runs differ in symbol, lengths skewed to short, '$' rare."""
    rng = np.random.default_rng(seed)
    sym = np.empty(nruns, dtype=np.uint8)
    step = rng.integers(1, 4, nruns).astype(np.uint8)          # next ACGT symbol differs from the previous
    sym[0] = 1
    acc = np.cumsum(step) % 4
    sym[:] = 1 + acc
    dollar = rng.random(nruns) < 0.01
    sym[dollar] = 0
    ln = np.minimum(31, 1 + rng.geometric(0.25, nruns)).astype(np.uint8)
    runs = (sym << 5) | ln
    return runs, int(ln.sum())


def vectors(R, n: int, totals, rng, nq=2000, npos=4000):
    pos = np.unique(np.concatenate([rng.integers(0, n, npos).astype(np.uint64),
                                    np.arange(max(0, n - 1500), n, dtype=np.uint64),
                                    np.arange(0, min(n, 200), dtype=np.uint64),
                                    np.array([np.iinfo(np.uint64).max], dtype=np.uint64)]))
    out = {"pos": pos}
    for c in b"$ACGT":
        out["occ_" + chr(c)] = R.get_occ_batch(np.full(pos.size, c, dtype=np.uint8), pos)
    inb = pos[pos < n]
    out["pos_in"] = inb
    out["char"] = R.get_char_batch(inb)
    out["f"] = np.array([ord(R.get_f(int(i))) for i in inb[:500]], dtype=np.uint8)
    for ci, c in enumerate(b"$ACGT"):
        cnt = np.unique(rng.integers(1, totals[ci] + 1, 600).astype(np.uint64)) if totals[ci] else np.zeros(0, np.uint64)
        out["cnt_" + chr(c)] = cnt
        out["occ_at_" + chr(c)] = R.get_occ_at_batch(np.full(cnt.size, c, dtype=np.uint8), cnt)
    return out


def read_set_golden(name: str, rng):
    """A real read-set BWT: `small` has a large reference tail region, `clean` an empty one."""
    case = common.make_case(name, *common.Datasets.SPECS[name][:3], min_len=73, keep_reads=True)
    bwt_path = os.path.join(HERE, name + ".bwt")
    common.write_case(case, bwt_path)
    R = oracle.RefIndex(bwt_path)
    n = case.num_symbols
    E = common.Emul(case)
    v = vectors(R, n, E.tot, rng)
    q = case.queries(2000, 31, seed=5)
    lo, up = R.find_interval_batch(q, 31)
    v.update(kmers=q, lower=lo, upper=up)
    # variable-length queries of the lengths in demo/test.in (37, 41, 46, 41, 100 x 5), cut from reads
    lens = [37, 41, 46, 41, 100, 100, 100, 100, 100]
    full = [r for r in case.reads if len(r) == 100]
    var = [full[(7 * i) % len(full)][: L] if L < 100 else full[(11 * i) % len(full)] for i, L in enumerate(lens)]
    vlo, vup = zip(*[R.find_interval(s) for s in var])
    v.update(var_queries=np.array(var), var_lower=np.array(vlo, dtype=np.uint64), var_upper=np.array(vup, dtype=np.uint64))
    rows = rng.integers(0, n, 2000).astype(np.uint64)
    lex, ps, ri, qs = R.locate_batch(rows, max_steps=300)
    v.update(rows=rows, lexrank=lex, pre_steps=ps, read_index=ri, post_steps=qs)
    # reads rebuilt by extractPrefix + extractPostfix for rows whose walks terminate
    ok = np.nonzero((ps != 0xFFFFFFFF) & (qs != 0xFFFFFFFF))[0][:64]
    v.update(read_rows=rows[ok], read_strings=np.array([R.extract_read(int(r)) for r in rows[ok]]))
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "x.bpi2")
        R.serialise(p)
        v["bpi2_sha256"] = np.array(hashlib.sha256(open(p, "rb").read()).hexdigest())
    v["sai"] = case.sai
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **v)
    print(name, ": n", n, "runs", case.runs.size, "hits", int((up >= lo).sum()), "walk cut-offs",
          int((ps == 0xFFFFFFFF).sum()), "tail rows", E.n - E.clamp_c)


def main():
    oracle.build(ref=True)
    assert oracle.have_ref(), "needs /root/reference"
    rng = np.random.default_rng(12345)

    for name in ("small", "clean"):
        read_set_golden(name, rng)

    # ---- deep_rand: depth-3 tree from seeded random runs ----
    seed, nruns = 777, (1 << 20) + 300_000
    runs, nsym = random_runs(seed, nruns)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "deep.bwt")
        import torch
        from readserver_b200 import datagen as dg
        dg.write_bwt(p, torch.from_numpy(runs), 1, nsym)
        R = oracle.RefIndex(p)
        tot = [int(((runs >> 5) == c).astype(np.uint64) @ (runs & 31).astype(np.uint64)) for c in range(5)]
        v = vectors(R, nsym, tot, rng)
        q = SYMS[rng.integers(1, 5, (2000, 12))]
        lo, up = R.find_interval_batch(q, 12)
        v.update(kmers=q, lower=lo, upper=up, seed=np.array(seed), nruns=np.array(nruns), nsym=np.array(nsym),
                 runs_sha256=np.array(hashlib.sha256(runs.tobytes()).hexdigest()))
        np.savez_compressed(os.path.join(HERE, "deep_rand.npz"), **v)
        print("deep_rand: n", nsym, "runs", nruns, "hits", int((up >= lo).sum()))


if __name__ == "__main__":
    main()
