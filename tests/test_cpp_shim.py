"""include/rs_gpu_bwt.hpp -- the C++ mirror of `class BWT` + query.h.  CPU: it compiles against the C ABI
(and, when /root/reference is there, against the reference's own bwt.h/query.h).  GPU: the program runs
and its answers equal the oracle's."""
import os
import subprocess

import numpy as np
import pytest

import oracle
from readserver_b200 import build as rs_build
from tests import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "shim_test.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "shim_test")
REF_INC = "/root/reference/include/bwt"


def _compile(extra=()):
    lib = rs_build.build()
    cmd = ["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), *extra, SRC, "-o", EXE,
           "-L", os.path.dirname(lib), "-lrs_b200", "-Wl,-rpath," + os.path.dirname(lib)]
    subprocess.run(cmd, check=True)
    return EXE


def test_shim_compiles_and_links():
    assert os.path.exists(_compile())


@pytest.mark.skipif(not os.path.isdir(REF_INC), reason="needs the reference headers")
def test_shim_compiles_against_reference_headers(tmp_path):
    """With the reference's bwt.h / query.h included first, GpuBWT derives from THEIR class BWT."""
    src = tmp_path / "with_ref.cpp"
    src.write_text('#include "bwt.h"\n#include "query.h"\n#include "rs_gpu_bwt.hpp"\n'
                   'uint64_t f(const BWT* p) { return p->getBWLen(); }\n'
                   'uint64_t g(rsb200::GpuBWT* b) { return f(b) + rsb200::findInterval(b, "ACGT").lower; }\n')
    subprocess.run(["g++", "-std=c++11", "-c", "-I", REF_INC, "-I", os.path.join(ROOT, "include"), str(src), "-o",
                    str(tmp_path / "with_ref.o")], check=True)


@pytest.mark.gpu
def test_shim_matches_oracle(tmp_path):
    exe = _compile()
    bwt = os.path.join(common.GOLDEN, "clean.bwt")
    z = np.load(os.path.join(common.GOLDEN, "clean.npz"))
    P = oracle.PortIndex(bwt)
    qs = [bytes(k).decode() for k in z["kmers"][:40]] + [str(s) for s in z["var_queries"]] + ["ACGTNACGT"]
    qf = tmp_path / "q.txt"
    qf.write_text("\n".join(qs) + "\n")
    sf = tmp_path / "clean.sai.u64"
    z["sai"].astype("<u8").tofile(str(sf))
    out = subprocess.run([exe, bwt, str(qf), str(sf)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[0] == f"bwlen {P.bwlen} pc " + " ".join(str(P.get_pc(c)) for c in "$ACGT")
    for i, (w, line) in enumerate(zip(qs, out[1:1 + len(qs)])):
        f = line.split()
        lo, up = P.find_interval(w)
        ok = all(c in "ACGT" for c in w)   # query() answers nothing for other bytes (query.cpp:90)
        assert (int(f[1]), int(f[2]), int(f[3]), int(f[4])) == (lo, up, lo, up), w
        if i < 16 or not ok:
            want = [P.extract_read(r) for r in range(lo, up + 1)] if ok and up >= lo else []
            if i < 16:
                assert int(f[5]) == len(want) and f[7:] == want
                assert int(f[6]) == int(w in want)
    tail = out[1 + len(qs)].split()
    n = P.bwlen
    assert tail[1] == P.get_f(0) + P.get_f(P.get_pc("C")) + P.get_f(n - 1)
    assert tail[3] == P.get_char(12345) and int(tail[5]) == P.get_occ_at("G", 777)
    # tileIntervalsBatch: get_tiles (service.cpp:232-246) with kmer 31, skip 4, one line per query
    for i, w in enumerate(qs):
        f = out[2 + len(qs) + i].split()
        tiles = [w[j:j + 31] for j in range(0, len(w) - 31 + 1, 5)] if len(w) >= 31 else []
        assert f[0] == "tiles" and int(f[1]) == len(tiles)
        got = [(int(f[2 + 2 * t]), int(f[3 + 2 * t])) for t in range(len(tiles))]
        assert got == [(1, 0) if any(c not in "ACGT" for c in t) else P.find_interval(t) for t in tiles], w
    # findGtReadsBatch: find_gt_reads over every tile (service.cpp:507-711), kmer 31, skip 7, intervals capped at 12
    from oracle import gt as ogt
    some = 0
    for i, w in enumerate(qs):
        f = out[2 + 2 * len(qs) + i].split()
        reads, undefined = ogt.gt_task_reads(P, w, len(w) // 2 + 1, 31, 7, 12)
        assert f[0] == "gt" and int(f[1]) == sum({"throw": 1, "hang": 2}[u] for u in undefined), w
        assert [int(x) for x in f[3:]] == reads and int(f[2]) == len(reads), w
        some += len(reads)
    assert some > 0
    assert out[2 + 3 * len(qs)] == "error 1"


@pytest.mark.gpu
def test_references_own_query_cpp_runs_over_gpubwt(tmp_path):
    """INTEGRATION.md section 2, proven: the reference's unmodified src/bwt/query.cpp -- findInterval, extractPrefix,
    extractPostfix, query, query_exactmatch -- compiled from /root/reference in the build container and linked against
    rsb200::GpuBWT (oracle/_ref/ref_query_over_gpubwt, built by oracle/Makefile), gives on the GPU what it gives over
    the reference's own RLEBWT (the oracle)."""
    import numpy as np
    import oracle
    from tests import common
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_query_over_gpubwt")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_query_over_gpubwt is built in the build container (needs /root/reference)")
    bwt = os.path.join(common.GOLDEN, "clean.bwt")
    z = np.load(os.path.join(common.GOLDEN, "clean.npz"))
    P = oracle.open_best(bwt) if oracle.have_ref() else oracle.PortIndex(bwt)
    Pp = oracle.PortIndex(bwt)
    gold = common.make_case("clean", *common.Datasets.SPECS["clean"][:3], min_len=73, keep_reads=True)
    full = [r for r in gold.reads if len(r) == 100]
    qs = [bytes(k).decode() for k in z["kmers"][:25]] + [full[1], full[2][:80], "ACGTNACGT", "GATTACA", full[3][20:60]]
    rows = [int(r) for r in z["read_rows"][:20]]
    (tmp_path / "q.txt").write_text("\n".join(qs) + "\n")
    (tmp_path / "r.txt").write_text("\n".join(str(r) for r in rows) + "\n")
    out = subprocess.run([exe, bwt, str(tmp_path / "q.txt"), str(tmp_path / "r.txt")], check=True, capture_output=True,
                         text=True).stdout.splitlines()
    assert len(out) == 3 * len(qs) + len(rows)
    for i, q in enumerate(qs):
        lo, up = P.find_interval(q)
        assert out[3 * i] == f"I {lo} {up}", q
        acgt = all(c in "ACGT" for c in q)
        reads = [Pp.extract_read(r) for r in range(lo, up + 1)] if acgt and up >= lo else []
        assert out[3 * i + 1] == f"X {int(acgt and q in reads)}", q
        assert out[3 * i + 2].split() == ["Q", str(len(reads))] + reads, q
    for j, r in enumerate(rows):
        assert out[3 * len(qs) + j] == f"R {Pp.extract_prefix(r)[0]}|{Pp.extract_postfix(r)[0]}"
    assert any(l.startswith("X 1") for l in out)
