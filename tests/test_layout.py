"""Host logic of the product on the CPU: the .bwt reader + block/directory builder (rs_build.cpp) and
the per-query rank/access/select code the kernels are compiled from (rs_ops.h, rs_swar.h), replayed on
host memory (tests/emul) and compared with the oracle.  No GPU needed."""
import hashlib
import os

import numpy as np
import pytest

import oracle
from tests import common
from tests.golden.make_golden import random_runs

U64MAX = np.iinfo(np.uint64).max


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_rank_access_select_match_oracle(name, datasets):
    case, P, E = datasets.case(name), datasets.port(name), common.Emul(datasets.case(name))
    assert E.status == 0 and E.q2_safe == 1 and E.depth == P.depth and E.vlast == P.vsum_size - 1
    assert E.pc == [P.get_pc(c) for c in "$ACGT"]
    assert E.init_occ == [P.get_occ(c, E.n - 1) for c in "$ACGT"]
    idx = np.concatenate([common.positions(E.n), np.array([U64MAX], dtype=np.uint64)])
    # "$ACGT" and bytes outside it: the reference's callers reject those except the /gt extension
    # loop; its getOcc then answers with '$' counts at 64-run bucket granularity (rs_occ_stray)
    for c in b"$ACGTN\x00z":
        s = np.full(idx.size, c, dtype=np.uint8)
        assert np.array_equal(E.occ(s, idx), P.get_occ_batch(s, idx)), chr(c)
    assert np.array_equal(E.char(idx[:-1]), P.get_char_batch(idx[:-1]))
    rng = np.random.default_rng(1)
    for ci, c in enumerate(b"$ACGT"):
        tot = E.tot[ci]
        cnt = np.arange(1, tot + 1, dtype=np.uint64) if tot <= 200_000 else rng.integers(1, tot + 1, 200_000).astype(np.uint64)
        s = np.full(cnt.size, c, dtype=np.uint8)
        assert np.array_equal(E.occ_at(s, cnt), P.get_occ_at_batch(s, cnt)), chr(c)


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_find_interval_and_walk_match_oracle(name, datasets):
    case, P, E = datasets.case(name), datasets.port(name), common.Emul(datasets.case(name))
    for k in (1, 2, 12, 31, 73):
        q = case.queries(1500, k, seed=40 + k)
        a, b = E.find_interval(q, k), P.find_interval_batch(q, k)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), k
    rows = np.random.default_rng(2).integers(0, E.n, 3000).astype(np.uint64)
    lex, st = E.prefix_walk(rows, 300)
    lex2, st2, _, _ = P.locate_batch(rows, postfix=False, max_steps=300)
    assert np.array_equal(lex, lex2) and np.array_equal(st, st2)
    ri, st = E.postfix_walk(rows, 300)
    _, _, ri2, st2 = P.locate_batch(rows, prefix=False, max_steps=300)
    assert np.array_equal(ri, ri2) and np.array_equal(st, st2)


def test_block_structure(datasets):
    """The 64-byte block as rs_layout.h documents it: start, 40-bit counters split into low words and
    high bytes, 36 run bytes (zero-padded in the last block), against a plain decode of the run bytes."""
    case = datasets.case("small")
    E = common.Emul(case)
    blocks = E.blocks()
    runs = case.runs
    assert E.nblocks == (runs.size + 35) // 36
    sym, ln = (runs >> 5).astype(np.int64), (runs & 31).astype(np.int64)
    csum = np.zeros((5, runs.size + 1), dtype=np.int64)
    for c in range(5):
        csum[c, 1:] = np.cumsum(np.where(sym == c, ln, 0))
    tot = np.concatenate([[0], np.cumsum(ln)])
    for b in (0, 1, E.nblocks // 2, E.nblocks - 1):
        raw = blocks[b]
        r0 = b * 36
        assert int(raw[0:8].view(np.uint64)[0]) == tot[r0]
        lo, hi = raw[8:24].view(np.uint32), raw[24:28]
        for c in range(4):
            assert int(lo[c]) + (int(hi[c]) << 32) == csum[c + 1, r0]
        body = runs[r0: min(r0 + 36, runs.size)]
        assert np.array_equal(raw[28: 28 + body.size], body) and not raw[28 + body.size:].any()


def test_without_tail_emulation_rank_is_exact(datasets):
    """emulate_tail = 0 must give the true rank/access everywhere (checked against the plain BWT)."""
    import torch
    from readserver_b200 import datagen as dg
    case = datasets.case("small")
    E = common.Emul(case)
    assert E.clamp_c < E.n  # the reference is wrong on this one's tail
    E.set_tail(False)
    bwt = dg.rle_decode(torch.from_numpy(case.runs)).numpy()
    idx = np.arange(E.n, dtype=np.uint64)
    for code, c in enumerate(b"$ACGT"):
        truth = np.cumsum(bwt == code).astype(np.uint64)
        assert np.array_equal(E.occ(np.full(idx.size, c, dtype=np.uint8), idx), truth)
    assert np.array_equal(E.char(idx), np.frombuffer(b"$ACGT", dtype=np.uint8)[bwt])
    q = case.queries(300, 31, seed=17)
    lo, up = E.find_interval(q, 31)
    cnt = np.where(up >= lo, up - lo + 1, 0)
    assert list(cnt) == [common.brute_count(case.reads, bytes(k).decode()) for k in q]


def test_layout_against_reference_golden():
    z = np.load(os.path.join(common.GOLDEN, "deep_rand.npz"))
    runs, nsym = random_runs(int(z["seed"]), int(z["nruns"]))
    case = common.Case("deep_rand", runs, 1, nsym, np.zeros(1, np.uint64), None)
    E = common.Emul(case)
    assert E.depth == 3
    for c in b"$ACGT":
        assert np.array_equal(E.occ(np.full(z["pos"].size, c, dtype=np.uint8), z["pos"]), z["occ_" + chr(c)])
        cnt = z["cnt_" + chr(c)]
        assert np.array_equal(E.occ_at(np.full(cnt.size, c, dtype=np.uint8), cnt), z["occ_at_" + chr(c)])
    assert np.array_equal(E.char(z["pos_in"]), z["char"])
    lo, up = E.find_interval(z["kmers"], z["kmers"].shape[1])
    assert np.array_equal(lo, z["lower"]) and np.array_equal(up, z["upper"])
    for name in ("small", "clean"):
        z = np.load(os.path.join(common.GOLDEN, name + ".npz"))
        E = common.Emul(path=os.path.join(common.GOLDEN, name + ".bwt"))
        lo, up = E.find_interval(z["kmers"], 31)
        assert np.array_equal(lo, z["lower"]) and np.array_equal(up, z["upper"])
        lex, st = E.prefix_walk(z["rows"], 300)
        assert np.array_equal(lex, z["lexrank"]) and np.array_equal(st, z["pre_steps"])


def test_ragged_and_degenerate_blocks():
    """Edge shapes of the block layout: runs of length 1 (more than three block boundaries per
    directory slot), maximal runs, a single run, '$'-only, block-exact run counts."""
    rng = np.random.default_rng(5)
    shapes = {
        "all_len1": ((rng.integers(0, 5, 5000).astype(np.uint8) << 5) | 1),
        "all_len31": ((rng.integers(1, 5, 3000).astype(np.uint8) << 5) | 31),
        "single_run": np.array([(2 << 5) | 7], dtype=np.uint8),
        "dollars": np.full(200, (0 << 5) | 31, dtype=np.uint8),
        "exact_36": ((rng.integers(0, 5, 36).astype(np.uint8) << 5) | rng.integers(1, 32, 36).astype(np.uint8)),
        "exact_72": ((rng.integers(0, 5, 72).astype(np.uint8) << 5) | rng.integers(1, 32, 72).astype(np.uint8)),
        "block_plus_1": ((rng.integers(0, 5, 37).astype(np.uint8) << 5) | rng.integers(1, 32, 37).astype(np.uint8)),
        "short_tail": ((rng.integers(0, 5, 97).astype(np.uint8) << 5) | rng.integers(1, 32, 97).astype(np.uint8)),
        "mixed": np.concatenate([(rng.integers(0, 5, 4000).astype(np.uint8) << 5) | 1,
                                 (rng.integers(0, 5, 4000).astype(np.uint8) << 5) | 31,
                                 (rng.integers(0, 5, 70000).astype(np.uint8) << 5) | rng.integers(1, 4, 70000).astype(np.uint8)]),
    }
    for name, runs in shapes.items():
        runs = np.ascontiguousarray(runs, dtype=np.uint8)
        n = int((runs & 31).astype(np.int64).sum())
        case = common.Case(name, runs, 1, n, np.zeros(1, np.uint64), None)
        E = common.Emul(case)
        assert E.status == 0, name
        P = oracle.PortIndex(runs=runs, num_strings=1, num_symbols=n)
        idx = np.concatenate([np.arange(n, dtype=np.uint64), np.array([U64MAX], dtype=np.uint64)])
        if not E.q2_safe:  # the reference itself would read out of bounds here; compare true ranks only
            E.set_tail(False)
            continue
        for c in b"$ACGT":
            s = np.full(idx.size, c, dtype=np.uint8)
            assert np.array_equal(E.occ(s, idx), P.get_occ_batch(s, idx)), (name, chr(c))
        assert np.array_equal(E.char(idx[:-1]), P.get_char_batch(idx[:-1])), name
        for ci, c in enumerate(b"$ACGT"):
            cnt = np.arange(1, E.tot[ci] + 1, dtype=np.uint64)
            assert np.array_equal(E.occ_at(np.full(cnt.size, c, np.uint8), cnt),
                                  P.get_occ_at_batch(np.full(cnt.size, c, np.uint8), cnt)), (name, chr(c))


def test_builder_rejects_bad_input(tmp_path):
    bad = np.array([(5 << 5) | 3], dtype=np.uint8)
    assert common.Emul(common.Case("b", bad, 1, 3, None, None)).status == 4
    zero = np.array([(1 << 5) | 0], dtype=np.uint8)
    assert common.Emul(common.Case("z", zero, 1, 0, None, None)).status == 4
    ok = np.array([(1 << 5) | 3], dtype=np.uint8)
    assert common.Emul(common.Case("m", ok, 1, 4, None, None)).status == 5  # header symbol count mismatch
    p = tmp_path / "bad.bwt"
    p.write_bytes(b"\x00" * 40)
    assert common.Emul(path=str(p)).status == 2
    p.write_bytes(b"\xca\xca" + b"\x00" * 10)
    assert common.Emul(path=str(p)).status == 3
    assert common.Emul(path=str(tmp_path / "missing.bwt")).status == 1


@pytest.mark.parametrize("sos,meta", [(1, False), (2, True), (1, True)])
def test_etl_sample_file_parser(sos, meta, datasets, tmp_path):
    """rs_parse_etl_samples: the merged `read / codes` file -> read-index-addressed CSR; reads with 'N' are not
    in the BWT and take no index; count/lane metadata bytes are dropped."""
    case = datasets.case("clean")
    rng = np.random.default_rng(3)
    n = case.num_strings
    cnt = rng.integers(1, 5, n)
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(cnt, out=off[1:])
    who = rng.integers(0, 200, int(off[-1]))
    codes = np.array([list(oracle.sample_code(int(i), sos)) for i in who], dtype=np.uint8)
    path = str(tmp_path / "merged.rlosorted.reads")
    common.write_etl_file(path, case.reads, off, codes, sos, meta, n_reads_every=97)
    rc, off2, codes2 = common.parse_etl(path, sos, meta)
    assert rc == 0 and np.array_equal(off2, off) and np.array_equal(codes2, codes.reshape(-1))
    # the value bytes of one read decode, with the oracle's restatement of service.cpp:829-843, to the same codes
    lines = open(path, "rb").read().split(b"\n")
    kept = [(lines[i], lines[i + 1]) for i in range(0, len(lines) - 1, 2) if b"N" not in lines[i]]
    for r in (0, 1, n // 2, n - 1):
        dec = oracle.decode_sample_value(kept[r][1], sos, meta)
        assert [d[0] for d in dec] == [bytes(c) for c in codes[int(off[r]):int(off[r + 1])]]
    assert common.parse_etl(str(tmp_path / "missing"), sos, meta)[0] == 1
    open(path, "ab").write(b"ACGT\n")          # a read without its value line
    assert common.parse_etl(path, sos, meta)[0] == 5


def test_fuzz_random_run_arrays_against_oracle():
    """Random small run arrays (every symbol, every run length, block-boundary sizes): the device code replayed on
    the host equals the oracle at every position; where the reference itself would read out of bounds (Q2) the
    exact mode is checked against a plain decode instead."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)
    @given(st.integers(1, 400), st.integers(0, 2**31 - 1), st.sampled_from([1, 2, 3, 31, 0]))
    def run(nruns, seed, maxlen):
        rng = np.random.default_rng(seed)
        ln = rng.integers(1, 32, nruns) if maxlen == 0 else rng.integers(1, maxlen + 1, nruns)
        runs = np.ascontiguousarray((rng.integers(0, 5, nruns) << 5 | ln).astype(np.uint8))
        n = int(ln.sum())
        E = common.Emul(common.Case("fuzz", runs, 1, n, np.zeros(1, np.uint64), None))
        assert E.status == 0
        idx = np.arange(n, dtype=np.uint64)
        plain = np.repeat(runs >> 5, runs & 31)
        if E.q2_safe:
            P = oracle.PortIndex(runs=runs, num_strings=1, num_symbols=n)
            for c in b"$ACGT":
                s = np.full(n, c, dtype=np.uint8)
                assert np.array_equal(E.occ(s, idx), P.get_occ_batch(s, idx))
            assert np.array_equal(E.char(idx), P.get_char_batch(idx))
        E.set_tail(False)
        for code, c in enumerate(b"$ACGT"):
            truth = np.cumsum(plain == code).astype(np.uint64)
            assert np.array_equal(E.occ(np.full(n, c, dtype=np.uint8), idx), truth)
            tot = int(truth[-1])
            if tot:
                want = np.nonzero(plain == code)[0].astype(np.uint64)
                assert np.array_equal(E.occ_at(np.full(tot, c, np.uint8), np.arange(1, tot + 1, dtype=np.uint64)), want)
        assert np.array_equal(E.char(idx), np.frombuffer(b"$ACGT", dtype=np.uint8)[plain])

    run()
