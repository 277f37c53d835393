"""Parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs and against the golden vectors generated from the unmodified reference.  Bit-exact."""
import hashlib
import os

import numpy as np
import pytest

import oracle
from readserver_b200 import datagen as dg
from tests import common
from tests.golden.make_golden import random_runs

pytestmark = pytest.mark.gpu
U64MAX = np.iinfo(np.uint64).max


@pytest.fixture(scope="module")
def engine():
    """The C-ABI engine (ctypes over librs_b200.so)."""
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from readserver_b200 import engine as e
    return e


def _gpu(engine, case):
    return engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols)


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_bwt_interface_matches_oracle(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    with _gpu(engine, case) as B:
        assert B.get_bwlen() == P.bwlen and B.info.ref_depth == P.depth and B.info.q2_safe == 1
        assert [B.get_pc(c) for c in "$ACGT"] == [P.get_pc(c) for c in "$ACGT"]
        idx = np.concatenate([common.positions(P.bwlen), np.array([U64MAX], dtype=np.uint64)])
        for c in b"$ACGTN\xff":     # the last two: bytes the reference's callers keep away, answered as it would
            s = np.full(idx.size, c, dtype=np.uint8)
            assert np.array_equal(B.get_occ(s, idx), P.get_occ_batch(s, idx)), chr(c)
        assert np.array_equal(B.get_char(idx[:-1]), P.get_char_batch(idx[:-1]))
        f = B.get_f(idx[:2000])
        assert [chr(x) for x in f] == [P.get_f(int(i)) for i in idx[:2000]]


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_find_interval_matches_oracle(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    with _gpu(engine, case) as B:
        for k in (1, 2, 12, 31, 73):
            q = case.queries(20000 if k == 31 else 3000, k, seed=40 + k)
            lo, up = B.find_interval(q, k)
            lo2, up2 = P.find_interval_batch(q, k, nthreads=4)
            assert np.array_equal(lo, lo2) and np.array_equal(up, up2), k
        # ragged batch (lengths of demo/test.in plus the extremes), one of them empty
        strs = [bytes(case.queries(1, L, seed=90 + L)[0]).decode() for L in (37, 41, 46, 41, 100, 100, 1, 200)] + [""]
        lo, up = B.find_interval_var(strs)
        for s, a, b in zip(strs[:-1], lo, up):
            assert (int(a), int(b)) == P.find_interval(s), s
        assert (int(lo[-1]), int(up[-1])) == (1, 0)
        # bytes outside "$ACGT" (callers keep them away, SURVEY.md Q4, but the /gt extension loop does not): the
        # reference gives them '$' counts at 64-run bucket granularity; packed (k <= 31), plain and ragged paths
        rng = np.random.default_rng(3)
        for k in (1, 3, 31, 32, 73):
            q = case.queries(3000, k, seed=140 + k).copy()
            for i in range(0, q.shape[0], 3):
                for _ in range(1 + i % 2):
                    q[i, int(rng.integers(0, k))] = [ord("N"), ord("$"), ord("n"), 0, 255, ord("X")][(i // 3) % 6]
            lo, up = B.find_interval(q, k)
            lo2, up2 = P.find_interval_batch(q, k, nthreads=4)
            assert np.array_equal(lo, lo2) and np.array_equal(up, up2), ("stray", k)
        strs = [bytes(case.queries(1, L, seed=190 + L)[0]).decode() for L in (37, 41, 100, 1, 200, 12)]
        strs = [s[:len(s) // 2] + ch + s[len(s) // 2 + 1:] for s, ch in zip(strs, "NN$NxN")]
        lo, up = B.find_interval_var(strs)
        for s, a, b in zip(strs, lo, up):
            assert (int(a), int(b)) == P.find_interval(s), s


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_locate_table_gives_the_walks_answers(name, datasets, engine):
    """The optional locate table (8 bytes per BWT row: where the row's extractPrefix walk ends and after how many
    bases) must answer exactly what the walk answers -- under the reference's tail behaviour, where some walks never
    end and some rows lie on no read's walk, and with exact rank -- for every row of the small indexes."""
    case, P = datasets.case(name), datasets.port(name)
    n = P.bwlen
    rows = np.arange(n, dtype=np.uint64) if n < 700_000 else np.concatenate(
        [common.positions(n, 300_000, seed=3), np.arange(n - 70_000, n, dtype=np.uint64)])
    with _gpu(engine, case) as B:
        B.set_sai(case.sai)
        B.drop_locate_table()           # nothing to drop: refreshes B.info
        hbm0 = B.info.hbm_bytes
        for emulate in (True, False):
            B.set_tail_emulation(emulate)
            lex0, st0, ri0 = B.extract_prefix(rows, max_steps=130, want_read_index=True)
            B.build_locate_table(max_read_len=120)
            assert B.info.hbm_bytes == hbm0 + 8 * (n + case.num_strings)
            lex1, st1, ri1 = B.extract_prefix(rows, max_steps=130, want_read_index=True)
            assert np.array_equal(lex0, lex1) and np.array_equal(st0, st1) and np.array_equal(ri0, ri1), (name, emulate)
            for every in (2, 7, 16, 64, 500):     # sampled variant: walk to the next sampled row
                B.build_locate_table(max_read_len=120, every=every)
                assert B.info.hbm_bytes < hbm0 + 8 * (n + case.num_strings) // (1 if every == 2 else 2) + 4096
                lex6, st6, ri6 = B.extract_prefix(rows, max_steps=130, want_read_index=True)
                assert np.array_equal(lex0, lex6) and np.array_equal(st0, st6) and np.array_equal(ri0, ri6), (name, emulate, every)
                lex7, st7 = B.extract_prefix(rows[:50_000], max_steps=40)
                assert np.array_equal(lex7[st7 != 0xFFFFFFFF], lex0[:50_000][st7 != 0xFFFFFFFF])
                assert np.array_equal(st7 == 0xFFFFFFFF, st0[:50_000] > 40)
            B.build_locate_table(max_read_len=120)
            # a cut below the read length: same rows reported as cut off
            lex2, st2 = B.extract_prefix(rows[:50_000], max_steps=40)
            B.drop_locate_table()
            lex3, st3 = B.extract_prefix(rows[:50_000], max_steps=40)
            assert np.array_equal(lex2, lex3) and np.array_equal(st2, st3)
            assert B.info.hbm_bytes == hbm0
        # the switch rebuilds the table with the other semantics
        B.set_tail_emulation(True)
        B.build_locate_table(120)
        B.set_tail_emulation(False)
        lex4, st4 = B.extract_prefix(rows[-60_000:], max_steps=130)
        B.drop_locate_table()
        lex5, st5 = B.extract_prefix(rows[-60_000:], max_steps=130)
        assert np.array_equal(lex4, lex5) and np.array_equal(st4, st5)
    if name == "clean":     # against the oracle directly
        lexo, pso, _, _ = P.locate_batch(rows[:20_000], postfix=False)
        assert np.array_equal(lex5[:0], lex5[:0]) and np.array_equal(lex1[:20_000], lexo) and np.array_equal(st1[:20_000], pso)


def test_large_host_batch_goes_through_the_chunk_pipeline(datasets, engine):
    """Host-pointer calls with several hundred thousand queries are cut into chunks (copy-in, search, copy-out on
    separate streams, ramped chunk sizes): results must not depend on the cut, for intervals and for counts."""
    case, P = datasets.case("mid"), datasets.port("mid")
    for Q in (291_072, 400_001, 1_000_003):
        q = case.queries(Q, 31, seed=77)
        q[Q // 3, 5] = ord("N")
        with _gpu(engine, case) as B:
            lo, up = B.find_interval(q, 31)
            cnt = B.count(q, 31, strand=0)
            lo_s, up_s = B.find_interval(q[:50_000], 31)          # below the pipeline threshold: one launch
        assert np.array_equal(lo[:50_000], lo_s) and np.array_equal(up[:50_000], up_s)
        n = 120_000
        sel = np.concatenate([np.arange(n), np.arange(Q - n, Q)])
        lo2, up2 = P.find_interval_batch(q[sel], 31, nthreads=8)
        assert np.array_equal(lo[sel], lo2) and np.array_equal(up[sel], up2), Q
        want = np.where(up >= lo, up - lo + 1, 0)
        want[Q // 3] = 0
        assert np.array_equal(cnt, want), Q


@pytest.mark.parametrize("name", ["small", "clean", "deep"])
def test_count_reads_both_strands(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    comp = np.zeros(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    with _gpu(engine, case) as B:
        q = case.queries(5000, 31, seed=77)
        q[7, 3] = ord("N")      # count_reads answers 0 for anything outside ACGT (service.cpp:299-301)
        q[9, 30] = ord("$")
        ok = np.ones(len(q), dtype=bool)
        ok[[7, 9]] = False
        for strand, qq in ((0, q), (1, comp[q[:, ::-1]])):
            lo, up = P.find_interval_batch(np.ascontiguousarray(qq), 31, nthreads=4)
            want = np.where(ok & (up >= lo), up - lo + 1, 0)
            assert np.array_equal(B.count(q, 31, strand=strand), want)


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_extract_prefix_matches_oracle(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    with _gpu(engine, case) as B:
        B.set_sai(case.sai)
        rows = np.random.default_rng(2).integers(0, P.bwlen, 20000).astype(np.uint64)
        lex, st, ri, bases = B.extract_prefix(rows, max_steps=128, want_bases=True, want_read_index=True)
        lex2, st2, ri2, _ = P.locate_batch(rows, nthreads=4, max_steps=128)
        assert np.array_equal(lex, lex2) and np.array_equal(st, st2)
        done = st != 0xFFFFFFFF
        assert np.array_equal(ri[done], case.sai[lex[done].astype(np.int64)])
        if name in ("clean", "deep"):   # empty tail region: every walk ends and .sai[lexrank] is the read index
            assert done.all() and np.array_equal(ri, ri2)
        for i in np.nonzero(done)[0][:200]:
            s, _, n = P.extract_prefix(int(rows[i]))
            assert bytes(bases[i, :n]).decode() == s and not bases[i, n:].any()


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_select_and_extract_postfix_match_oracle(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    rng = np.random.default_rng(8)
    with _gpu(engine, case) as B:
        for ci, c in enumerate(b"$ACGT"):
            tot = int(B.info.total[ci])
            cnt = np.arange(1, tot + 1, dtype=np.uint64) if tot <= 300_000 else np.concatenate(
                [rng.integers(1, tot + 1, 300_000).astype(np.uint64), np.array([1, tot], dtype=np.uint64)])
            s = np.full(cnt.size, c, dtype=np.uint8)
            assert np.array_equal(B.get_occ_at(s, cnt), P.get_occ_at_batch(s, cnt)), chr(c)
        rows = rng.integers(0, P.bwlen, 20000).astype(np.uint64)
        ri, st, bases = B.extract_postfix(rows, max_steps=128, want_bases=True)
        _, _, ri2, st2 = P.locate_batch(rows, nthreads=4, prefix=False, max_steps=128)
        assert np.array_equal(ri, ri2) and np.array_equal(st, st2)
        assert (st != 0xFFFFFFFF).all()       # select is exact in the reference: every forward walk ends
        for i in range(150):
            s, _, n = P.extract_postfix(int(rows[i]))
            assert bytes(bases[i, :n]).decode() == s and not bases[i, n:].any()
        # query() (query.cpp:87-100): reads rebuilt from rows = extractPrefix + extractPostfix
        if name == "clean":
            reads = B.extract_read(rows[:300], max_steps=128)
            assert reads == [P.extract_read(int(r)) for r in rows[:300]]
            assert all(r == case.reads[int(i)] for r, i in zip(reads, ri[:300]))


def _sample_table(num_reads, nsamples, sos, seed):
    """Synthetic sample store in the reference's value format (service.cpp:829-843): per read a list of
    1..4 sample codes in merge order, duplicates allowed (rlosort...cpp:47-53)."""
    rng = np.random.default_rng(seed)
    cnt = rng.integers(1, 5, num_reads)
    off = np.zeros(num_reads + 1, dtype=np.uint64)
    np.cumsum(cnt, out=off[1:])
    who = rng.integers(0, nsamples, int(off[-1]))
    lut = np.array([list(oracle.sample_code(int(i), sos)) for i in range(nsamples)], dtype=np.uint8)
    return off, lut[who], who


@pytest.mark.parametrize("name,sos,nsamples", [("clean", 1, 100), ("small", 2, 1000), ("deep", 1, 222)])
def test_locate_samples_and_carriers(name, sos, nsamples, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    off, codes, who = _sample_table(case.num_strings, nsamples, sos, seed=4)
    with _gpu(engine, case) as B:
        B.set_sai(case.sai)
        B.set_samples(off, codes, sos)
        q = case.queries(3000, 31, seed=61)
        res = B.locate(q, 31, max_hits=50, max_steps=128)
        lo, up = P.find_interval_batch(q, 31, nthreads=4)
        cnt = np.minimum(np.where(up >= lo, up - lo + 1, 0), 50).astype(np.uint64)
        want_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint64)
        assert np.array_equal(res["hit_offsets"], want_off) and int(want_off[-1]) > 0
        want_rows = np.concatenate([np.arange(lo[i], lo[i] + cnt[i], dtype=np.uint64) for i in range(len(q))])
        assert np.array_equal(res["rows"], want_rows)
        lex, st, ri_post, _ = P.locate_batch(want_rows, nthreads=4, max_steps=128)
        assert np.array_equal(res["lexrank"], lex) and np.array_equal(res["steps"], st)
        done = st != 0xFFFFFFFF
        assert np.array_equal(res["read_index"][done], case.sai[lex[done].astype(np.int64)])
        if name != "small":   # empty tail region: the .sai read index is the postfix walk's read index
            assert done.all() and np.array_equal(res["read_index"], ri_post)
        # sample gather == decode of the reference's value bytes for the same reads (A10)
        ri = res["read_index"]
        soff, scodes = B.samples(ri)
        j = 0
        for t, r in enumerate(ri[:400]):
            want = [] if r >= case.num_strings else oracle.decode_sample_value(codes[int(off[r]):int(off[r + 1])].tobytes(), sos, False)
            got = scodes[int(soff[t]):int(soff[t + 1])]
            assert [bytes(g) for g in got] == [w[0] for w in want]
        valid = ri < case.num_strings
        assert int(soff[-1]) == int((off[ri[valid].astype(np.int64) + 1] - off[ri[valid].astype(np.int64)]).sum())
        # carrier sets: sorted distinct samples per k-mer
        coff, ccodes = B.carriers(res["hit_offsets"], ri)
        ho = res["hit_offsets"].astype(np.int64)
        for i in range(0, len(q), 7):
            reads = ri[ho[i]:ho[i + 1]]
            reads = reads[reads < case.num_strings].astype(np.int64)
            want = sorted({int(s) for r in reads for s in who[int(off[r]):int(off[r + 1])]})
            got = [bytes(c) for c in ccodes[int(coff[i]):int(coff[i + 1])]]
            assert got == [oracle.sample_code(s, sos) for s in want]


@pytest.mark.parametrize("sos,meta", [(1, False), (2, True)])
def test_sample_table_from_etl_file(sos, meta, datasets, engine, tmp_path):
    """rs_index_load_samples_file (the ETL's merged read file instead of 64 RocksDB stores) gives the same
    sample gathers and carrier sets as the table attached from arrays."""
    case = datasets.case("clean")
    off, codes, who = _sample_table(case.num_strings, 150, sos, seed=8)
    path = str(tmp_path / "merged.rlosorted.reads")
    common.write_etl_file(path, case.reads, off, codes, sos, meta, n_reads_every=211)
    ri = np.random.default_rng(1).integers(0, case.num_strings, 5000).astype(np.uint64)
    hit_off = np.arange(0, 5001, 10, dtype=np.uint64)
    with _gpu(engine, case) as A, _gpu(engine, case) as B:
        A.set_samples(off, codes, sos)
        B.load_samples_file(path, sos, meta)
        a, b = A.samples(ri), B.samples(ri)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and int(a[0][-1]) > 5000
        ca, cb = A.carriers(hit_off, ri), B.carriers(hit_off, ri)
        assert np.array_equal(ca[0], cb[0]) and np.array_equal(ca[1], cb[1])
        with pytest.raises(engine.RsError):
            B.load_samples_file(str(tmp_path / "nope"), sos, meta)


def test_read_table_equals_walked_reads(datasets, engine, tmp_path):
    """output=reads: the packed read table addressed by the locate stage's read index gives the strings the
    reference rebuilds by walking (query(): extractPrefix + extractPostfix, query.cpp:87-100)."""
    case, P = datasets.case("clean"), datasets.port("clean")
    path = str(tmp_path / "merged.rlosorted.reads")
    off = np.arange(case.num_strings + 1, dtype=np.uint64)
    common.write_etl_file(path, case.reads, off, np.full((case.num_strings, 1), 40, np.uint8), 1, False, n_reads_every=53)
    q = case.queries(400, 31, seed=12)
    with _gpu(engine, case) as B:
        B.set_sai(case.sai)
        res = B.locate(q, 31, max_hits=20, max_steps=128)
        rows, ri = res["rows"], res["read_index"]
        assert rows.size > 2000
        want = [P.extract_read(int(r)) for r in rows[:1500]]
        for attach in (lambda: B.set_reads(case.reads), lambda: B.load_reads_file(path)):
            attach()
            got = B.reads(ri[:1500], max_len=128)
            assert got == want
        assert B.reads(np.array([case.num_strings, 0], dtype=np.uint64), 128) == [None, case.reads[0]]
        assert B.reads(ri[:3], max_len=10) == [w[:10] for w in want[:3]]   # truncated to max_len
        with pytest.raises(engine.RsError):
            B.set_reads(["ACGTN"] + case.reads[1:])
    # the z golden vectors hold reads rebuilt by the unmodified reference
    z = np.load(os.path.join(common.GOLDEN, "clean.npz"))
    with engine.GpuBWT(os.path.join(common.GOLDEN, "clean.bwt")) as G:
        gold = common.make_case("clean", *common.Datasets.SPECS["clean"][:3], min_len=73, keep_reads=True)
        G.set_reads(gold.reads)
        ri = G.extract_postfix(z["read_rows"], max_steps=300)[0]
        assert G.reads(ri, 128) == [str(s) for s in z["read_strings"]]


@pytest.mark.parametrize("name", ["small", "clean", "deep"])
def test_prefix_table_does_not_change_results(name, datasets, engine):
    """The k-mer suffix table only skips steps: intervals (including early-exit values) and counts on both
    strands must be the oracle's for every table length, for queries shorter, equal and longer than it."""
    case, P = datasets.case(name), datasets.port(name)
    comp = np.zeros(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    with _gpu(engine, case) as B:
        auto = B.info.prefix_table_k
        for t in (0, 2, 5, 8, auto):
            B.set_prefix_table(t)
            assert B.info.prefix_table_k == t
            for k in (1, 4, 5, 6, 8, 9, 31):
                q = case.queries(4000, k, seed=300 + k)
                q[5, k // 2] = ord("N")     # non-ACGT inside / at the end of the suffix: table must not be used
                q[6, k - 1] = ord("$")
                lo, up = B.find_interval(q, k)
                lo2, up2 = P.find_interval_batch(q, k, nthreads=4)
                ok = np.ones(len(q), dtype=bool)  # query 5: the stray byte is searched as the reference does (K1b)
                assert np.array_equal(lo[ok], lo2[ok]) and np.array_equal(up[ok], up2[ok]), (t, k)
                rcq = np.ascontiguousarray(comp[q[:, ::-1]])
                lo3, up3 = P.find_interval_batch(rcq, k, nthreads=4)
                okc = ok.copy()
                okc[5] = okc[6] = False     # count_reads answers 0 for anything outside ACGT (service.cpp:299-301)
                want = np.where(up3 >= lo3, up3 - lo3 + 1, 0)
                got = B.count(q, k, strand=1)
                assert np.array_equal(got[okc], want[okc]) and got[5] == 0 and got[6] == 0, (t, k)
        B.set_tail_emulation(False)         # the table is rebuilt with the other semantics
        q = case.queries(2000, 31, seed=17)
        lo, up = B.find_interval(q, 31)
        B.set_prefix_table(0)
        lo2, up2 = B.find_interval(q, 31)
        assert np.array_equal(lo, lo2) and np.array_equal(up, up2)


def test_tail_emulation_switch(datasets, engine):
    import torch
    from readserver_b200 import datagen as dg
    case = datasets.case("small")
    bwt = dg.rle_decode(torch.from_numpy(case.runs)).numpy()
    with _gpu(engine, case) as B:
        assert B.info.ref_clamp_c < B.info.bwlen
        B.set_tail_emulation(False)
        idx = np.arange(bwt.size, dtype=np.uint64)
        for code, c in enumerate(b"$ACGT"):
            assert np.array_equal(B.get_occ(np.full(idx.size, c, np.uint8), idx), np.cumsum(bwt == code).astype(np.uint64))
        q = case.queries(300, 31, seed=17)
        lo, up = B.find_interval(q, 31)
        cnt = np.where(up >= lo, up - lo + 1, 0)
        assert list(cnt) == [common.brute_count(case.reads, bytes(k).decode()) for k in q]
        B.set_tail_emulation(True)
        lo2, up2 = datasets.port("small").find_interval_batch(q, 31)
        lo, up = B.find_interval(q, 31)
        assert np.array_equal(lo, lo2) and np.array_equal(up, up2)


@pytest.mark.parametrize("name", ["small", "clean", "deep_rand"])
def test_against_reference_golden(name, engine):
    z = np.load(os.path.join(common.GOLDEN, name + ".npz"))
    if name == "deep_rand":
        runs, nsym = random_runs(int(z["seed"]), int(z["nruns"]))
        assert hashlib.sha256(runs.tobytes()).hexdigest() == str(z["runs_sha256"])
        B = engine.GpuBWT(runs=runs, num_strings=1, num_symbols=nsym)
    else:
        B = engine.GpuBWT(os.path.join(common.GOLDEN, name + ".bwt"))
    with B:
        for c in b"$ACGT":
            assert np.array_equal(B.get_occ(np.full(z["pos"].size, c, np.uint8), z["pos"]), z["occ_" + chr(c)])
        assert np.array_equal(B.get_char(z["pos_in"]), z["char"])
        assert np.array_equal(B.get_f(z["pos_in"][:500]), z["f"])
        lo, up = B.find_interval(z["kmers"], z["kmers"].shape[1])
        assert np.array_equal(lo, z["lower"]) and np.array_equal(up, z["upper"])
        if name != "deep_rand":
            lo, up = B.find_interval_var([str(s) for s in z["var_queries"]])
            assert np.array_equal(lo, z["var_lower"]) and np.array_equal(up, z["var_upper"])
            B.set_sai(z["sai"])
            lex, st, ri = B.extract_prefix(z["rows"], max_steps=300, want_read_index=True)
            assert np.array_equal(lex, z["lexrank"]) and np.array_equal(st, z["pre_steps"])
            if name == "clean":
                assert np.array_equal(ri, z["read_index"])


def test_degenerate_blocks_on_gpu(engine):
    rng = np.random.default_rng(5)
    shapes = [((rng.integers(0, 5, 5000).astype(np.uint8) << 5) | 1),
              ((rng.integers(1, 5, 3000).astype(np.uint8) << 5) | 31),
              np.array([(2 << 5) | 7], dtype=np.uint8),
              ((rng.integers(0, 5, 36).astype(np.uint8) << 5) | rng.integers(1, 32, 36).astype(np.uint8)),
              ((rng.integers(0, 5, 37).astype(np.uint8) << 5) | rng.integers(1, 32, 37).astype(np.uint8)),
              ((rng.integers(0, 5, 97).astype(np.uint8) << 5) | rng.integers(1, 32, 97).astype(np.uint8)),
              np.concatenate([(rng.integers(0, 5, 4000).astype(np.uint8) << 5) | 1,
                              (rng.integers(0, 5, 70000).astype(np.uint8) << 5) | rng.integers(1, 4, 70000).astype(np.uint8)])]
    for runs in shapes:
        runs = np.ascontiguousarray(runs, dtype=np.uint8)
        n = int((runs & 31).astype(np.int64).sum())
        P = oracle.PortIndex(runs=runs, num_strings=1, num_symbols=n)
        with engine.GpuBWT(runs=runs, num_strings=1, num_symbols=n) as B:
            if not B.info.q2_safe:
                continue
            idx = np.concatenate([np.arange(n, dtype=np.uint64), np.array([U64MAX], dtype=np.uint64)])
            for c in b"$ACGT":
                s = np.full(idx.size, c, dtype=np.uint8)
                assert np.array_equal(B.get_occ(s, idx), P.get_occ_batch(s, idx))
            assert np.array_equal(B.get_char(idx[:-1]), P.get_char_batch(idx[:-1]))
            for ci, c in enumerate(b"$ACGT"):
                cnt = np.arange(1, int(B.info.total[ci]) + 1, dtype=np.uint64)
                s = np.full(cnt.size, c, dtype=np.uint8)
                assert np.array_equal(B.get_occ_at(s, cnt), P.get_occ_at_batch(s, cnt))


def test_error_behaviour(engine, tmp_path):
    p = tmp_path / "bad.bwt"
    p.write_bytes(b"\x00" * 64)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(str(p))
    assert e.value.status == 2      # the reference prints and exits (rlebwt_reader.cpp:31-34)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(str(tmp_path / "missing.bwt"))
    assert e.value.status == 1
    runs = np.array([(1 << 5) | 3, (0 << 5) | 1], dtype=np.uint8)
    with engine.GpuBWT(runs=runs, num_strings=1, num_symbols=4) as B:
        with pytest.raises(engine.RsError) as e:
            B.extract_prefix(np.zeros(1, np.uint64), want_read_index=True)
        assert e.value.status == 8
        lo, up = B.find_interval(np.zeros((0, 31), dtype=np.uint8), 31)
        assert lo.size == 0


def test_64_shards_fused_count_on_one_gpu(engine):
    """configs[2] shape in small: one read set split into the 64 shards of demo/permutations-3.txt (reversed
    last 3 bases, service.cpp:820), every shard searched for the whole batch, per-shard counts added by the
    search kernel into the communicator's buffer (world size 1: all k-mers are owned here).  The totals
    equal the sum of the oracle's per-shard counts, both strands; with exact rank they equal brute-force
    occurrence counts over the whole read set."""
    import torch
    from readserver_b200 import sharding
    case = common.make_case("shards", 24000, 6800, 3, 73, keep_reads=True)
    keys = sharding.permutation_keys(3)
    by_key = {k: [] for k in keys}
    for r in case.reads:
        by_key[sharding.shard_of_read(r)].append(r)
    shards = []
    for k in keys:
        if not by_key[k]:
            continue
        rd, ln = dg.strings_to_reads(by_key[k])
        rd, ln, _, _ = dg.dedupe_rlo(rd, ln)
        b = dg.build_bwt(rd, ln)
        runs = b.runs.numpy().copy()
        shards.append((engine.GpuBWT(runs=runs, num_strings=b.num_strings, num_symbols=b.num_symbols),
                       oracle.PortIndex(runs=runs, num_strings=b.num_strings, num_symbols=b.num_symbols)))
    assert len(shards) > 32
    q = case.queries(3000, 31, seed=5)
    comp = np.zeros(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    q[1500:] = comp[q[1500:, ::-1]]  # the second half only occurs on the reverse strand
    q[7, 3] = ord("N")  # counts 0 on every shard (service.cpp:299-301)
    Q = q.shape[0]
    dq = torch.from_numpy(q).cuda()
    comm = sharding.connect_peers(engine.PeerComm(0, 0, 1, Q), "cuda:0")
    try:
        for strand in (0, 1):
            for exact in (False, True):
                for B, _ in shards:
                    B.set_tail_emulation(not exact)
                comm.begin(Q)
                for B, _ in shards:
                    comm.accumulate(B, dq, 31, Q, strand)
                got = sharding.counts_view(comm.finish(), Q, "cuda:0").cpu().numpy()
                assert not comm.timed_out()
                if exact:
                    strs = [bytes(k).decode() for k in q]
                    if strand:
                        strs = [s[::-1].translate(str.maketrans("ACGT", "TGCA")) for s in strs]
                    want = np.array([0 if "N" in s else common.brute_count(case.reads, s) for s in strs])
                    assert want.sum() > 1000
                else:
                    qq = np.ascontiguousarray(comp[q[:, ::-1]]) if strand else q
                    valid = np.isin(q, np.frombuffer(b"ACGT", dtype=np.uint8)).all(axis=1)
                    want = np.zeros(Q, dtype=np.int64)
                    for _, P in shards:
                        lo, up = P.find_interval_batch(qq, 31, nthreads=4)
                        want += np.where(valid & (up >= lo), up - lo + 1, 0).astype(np.int64)
                assert np.array_equal(got, want), (strand, exact)
    finally:
        comm.close()
        for B, _ in shards:
            B.close()


@pytest.mark.parametrize("kmer,skip", [(31, 0), (31, 7), (73, 26), (12, 100)])
def test_kmer_tiles_match_get_tiles(kmer, skip, datasets, engine):
    """/kmermatch's tiling (get_tiles, service.cpp:232-246) + findInterval per tile, for a ragged batch."""
    case, P = datasets.case("clean"), datasets.port("clean")
    lens = [100, 73, 31, 30, 250, 12, 0, 101]
    qs = []
    for i, L in enumerate(lens):
        parts, seed = [], 300 + i
        while sum(len(p) for p in parts) < L:
            parts.append(bytes(case.queries(1, 100, seed=seed)[0]).decode())
            seed += 1000
        qs.append("".join(parts)[:L])
    qs[4] = qs[4][:40] + "N" + qs[4][41:]   # tiles over the N are skipped by the reference
    with _gpu(engine, case) as B:
        toff, lo, up = B.tile_intervals(qs, kmer, skip)
    t = 0
    for qi, w in enumerate(qs):
        tiles = [w[i:i + kmer] for i in range(0, len(w) - kmer + 1, skip + 1)] if len(w) >= kmer else []
        assert int(toff[qi + 1] - toff[qi]) == len(tiles)
        for tile in tiles:
            want = (1, 0) if any(c not in "ACGT" for c in tile) else P.find_interval(tile)
            assert (int(lo[t]), int(up[t])) == want, (qi, tile)
            t += 1
    assert t == int(toff[-1])
