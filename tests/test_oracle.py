"""The CPU oracle (oracle/rs_oracle.c, a plain-C restatement of the reference) is pinned here:
against golden vectors generated from the unmodified reference (tests/golden/make_golden.py), against
the unmodified reference itself when oracle/_ref/libref_bwt.so is present, and against brute-force
truth where the reference is exact."""
import hashlib
import os

import numpy as np
import pytest

import oracle
from tests import common
from tests.golden.make_golden import random_runs

G = common.GOLDEN
U64MAX = np.iinfo(np.uint64).max


def _golden_index(name):
    z = np.load(os.path.join(G, name + ".npz"))
    if name == "deep_rand":
        runs, nsym = random_runs(int(z["seed"]), int(z["nruns"]))
        assert hashlib.sha256(runs.tobytes()).hexdigest() == str(z["runs_sha256"])
        return z, oracle.PortIndex(runs=runs, num_strings=1, num_symbols=nsym)
    return z, oracle.PortIndex(os.path.join(G, name + ".bwt"))


@pytest.mark.parametrize("name", ["small", "clean", "deep_rand"])
def test_port_matches_reference_golden(name):
    z, P = _golden_index(name)
    pos = z["pos"]
    for c in b"$ACGT":
        got = P.get_occ_batch(np.full(pos.size, c, dtype=np.uint8), pos)
        assert np.array_equal(got, z["occ_" + chr(c)]), f"getOcc({chr(c)})"
    assert np.array_equal(P.get_char_batch(z["pos_in"]), z["char"])
    assert [ord(P.get_f(int(i))) for i in z["pos_in"][:500]] == list(z["f"])
    for c in b"$ACGT":
        cnt = z["cnt_" + chr(c)]
        assert np.array_equal(P.get_occ_at_batch(np.full(cnt.size, c, dtype=np.uint8), cnt), z["occ_at_" + chr(c)])
    k = z["kmers"].shape[1]
    lo, up = P.find_interval_batch(z["kmers"], k)
    assert np.array_equal(lo, z["lower"]) and np.array_equal(up, z["upper"])


@pytest.mark.parametrize("name", ["small", "clean"])
def test_port_locate_and_files_match_reference_golden(name, tmp_path):
    z, P = _golden_index(name)
    lex, ps, ri, qs = P.locate_batch(z["rows"], max_steps=300)
    assert np.array_equal(lex, z["lexrank"]) and np.array_equal(ps, z["pre_steps"])
    assert np.array_equal(ri, z["read_index"]) and np.array_equal(qs, z["post_steps"])
    for row, s in zip(z["read_rows"], z["read_strings"]):
        assert P.extract_read(int(row)) == str(s)
    for s, lo, up in zip(z["var_queries"], z["var_lower"], z["var_upper"]):
        assert P.find_interval(str(s)) == (int(lo), int(up))
    p = str(tmp_path / "x.bpi2")
    P.write_bpi2(p)  # byte-identical to the file the reference's serialiseFMIndex wrote
    assert hashlib.sha256(open(p, "rb").read()).hexdigest() == str(z["bpi2_sha256"])
    P2 = oracle.PortIndex(os.path.join(G, name + ".bwt"))
    assert np.array_equal(P2.get_occ_batch(np.full(50, ord("G"), dtype=np.uint8), z["pos"][:50]), z["occ_G"][:50])


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libref_bwt.so not built (needs /root/reference)")
@pytest.mark.parametrize("name", ["tiny", "small", "clean"])
def test_port_vs_live_reference_every_position(name, datasets, tmp_path):
    case = datasets.case(name)
    path = common.write_case(case, str(tmp_path / "t.bwt"))
    P, R = datasets.port(name), oracle.RefIndex(path)
    n = case.num_symbols
    idx = np.concatenate([np.arange(n, dtype=np.uint64), np.array([U64MAX], dtype=np.uint64)])
    for c in b"$ACGTN":
        s = np.full(idx.size, c, dtype=np.uint8)
        assert np.array_equal(P.get_occ_batch(s, idx), R.get_occ_batch(s, idx))
    assert np.array_equal(P.get_char_batch(idx[:-1]), R.get_char_batch(idx[:-1]))
    E = common.Emul(case)
    for ci, c in enumerate(b"$ACGT"):
        cnt = np.arange(1, E.tot[ci] + 1, dtype=np.uint64)
        s = np.full(cnt.size, c, dtype=np.uint8)
        assert np.array_equal(P.get_occ_at_batch(s, cnt), R.get_occ_at_batch(s, cnt))
    q = case.queries(3000, 31, seed=9)
    for nt in (1, 3):
        a, b = P.find_interval_batch(q, 31, nthreads=nt), R.find_interval_batch(q, 31, nthreads=nt)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    rows = np.random.default_rng(3).integers(0, n, 1500).astype(np.uint64)
    for x, y in zip(P.locate_batch(rows, max_steps=300), R.locate_batch(rows, max_steps=300)):
        assert np.array_equal(x, y)
    # the reference's own .bpi2 writer (src/util/index_rlebwt.cpp) and loader
    R.serialise(path + ".bpi2")
    p2 = str(tmp_path / "port.bpi2")
    P.write_bpi2(p2)
    assert open(p2, "rb").read() == open(path + ".bpi2", "rb").read()
    P3 = oracle.PortIndex(path, use_bpi2=True)
    assert np.array_equal(P3.get_occ_batch(np.full(2000, ord("T"), np.uint8), idx[-2001:-1]),
                          R.get_occ_batch(np.full(2000, ord("T"), np.uint8), idx[-2001:-1]))


def test_oracle_equals_truth_where_reference_is_exact(datasets):
    """`clean` has an empty tail region, so the reference must agree with brute force."""
    case, P = datasets.case("clean"), datasets.port("clean")
    E = common.Emul(case)
    assert E.clamp_c == E.n
    q = case.queries(300, 31, seed=17)
    lo, up = P.find_interval_batch(q, 31)
    cnt = np.where(up >= lo, up - lo + 1, 0)
    truth = [common.brute_count(case.reads, bytes(k).decode()) for k in q]
    assert list(cnt) == truth and sum(truth) > 1000
    # locate: postfix terminal = input-order read index; prefix terminal = lexicographic rank; .sai links them
    hit = np.nonzero(up >= lo)[0][:60]
    rows = np.concatenate([np.arange(lo[i], up[i] + 1, dtype=np.uint64) for i in hit])
    lex, ps, ri, qs = P.locate_batch(rows, max_steps=300)
    assert np.array_equal(case.sai[lex.astype(np.int64)], ri)
    j = 0
    for i in hit:
        kmer = bytes(q[i]).decode()
        for _ in range(int(up[i] - lo[i] + 1)):
            read = case.reads[int(ri[j])]
            assert read[int(ps[j]): int(ps[j]) + 31] == kmer and len(read) == int(ps[j]) + int(qs[j])
            assert P.extract_read(int(rows[j])) == read
            j += 1


def test_bwt_generator_equals_reference_builder(datasets, tmp_path):
    """datagen's suffix-sort BWT is byte-identical to the reference pipeline's builder (SGA's bcr.c)."""
    if not oracle.have_bcr():
        pytest.skip("oracle/_ref/libref_bcr.so not built")
    for name in ("tiny", "small"):
        case = datasets.case(name)
        a = common.write_case(case, str(tmp_path / (name + ".bwt")))
        oracle.ref_bcr_build(case.reads, a + ".bcr")
        assert open(a, "rb").read() == open(a + ".bcr", "rb").read()


def test_sample_value_codec():
    # create_samples_hash.cpp:31-51 and service.cpp:816-845
    assert oracle.sample_code(0, 1) == b'"' and oracle.sample_code(221, 1) == bytes([33 + 222])
    assert oracle.sample_code(222, 2) == bytes([33 + 1, 33 + 0])
    v = oracle.sample_code(5, 2) + bytes([33 + 3, 33 + 7]) + oracle.sample_code(700, 2) + bytes([33 + 1, 33 + 2])
    assert oracle.decode_sample_value(v, 2, True) == [(oracle.sample_code(5, 2), 3, 7), (oracle.sample_code(700, 2), 1, 2)]
    assert oracle.decode_sample_value(b"ab", 1, False) == [(b"a", 0, 0), (b"b", 0, 0)]
    assert oracle.decode_sample_value(b"", 1, False) == []
