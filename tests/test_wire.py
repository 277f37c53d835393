"""include/rs_wire.hpp and include/rs_service.hpp: the worker's protobuf messages and the batched count path.
CPU: the C++ codec's bytes equal google.protobuf's for the schema of readserver.proto.  GPU: a batch of
serialised Requests through CountBatcher yields the Replies count_reads (service.cpp:279-315) would send."""
import os
import subprocess

import numpy as np
import pytest

from readserver_b200 import build as rs_build
from tests import common, proto_schema

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")


def _wire_exe():
    exe = os.path.join(CPP, "wire_test")
    subprocess.run(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(CPP, "wire_test.cpp"), "-o", exe], check=True)
    return exe


def _run(exe, lines):
    out = subprocess.run([exe], input="\n".join(lines) + "\n", check=True, capture_output=True, text=True).stdout
    return out.splitlines()


def test_reply_bytes_equal_protobuf():
    M = proto_schema.build()
    exe = _wire_exe()
    cmds, want = [], []
    for rt, q, rev, cnt in [(1, "ACGT", 0, 5), (2, "ACGTACGTAC" * 10, 1, 0), (1, "", 0, 300), (1, "NNN", 1, 2**31 + 7),
                            (2, "T", 0, 2**32 + 129), (1, "ACGT", 1, 2**63)]:
        cmds.append(f"count\t{rt}\t{q}\t{rev}\t{cnt}")
        want.append(proto_schema.count_reply(M, rt, q, bool(rev), cnt).hex())
    # reads
    for rev, reads in [(0, ["ACGTT", "GGA"]), (1, ["A" * 300]), (0, [])]:
        r = M["Reply"]()
        r.rt, r.t, r.q = 2, 2, "ACG"
        r.r.SetInParent()
        for rd in reads:
            (r.r.revcomp_matches if rev else r.r.forward_matches).add().r = rd
        cmds.append(f"reads\t2\tACG\t{rev}\t{','.join(reads)}")
        want.append(r.SerializeToString().hex())
    # all
    hits = [("ACGTA", [("s1", 3, 7), ("sample_two", -1, 0)]), ("GGGT", [])]
    for rev in (0, 1):
        r = M["Reply"]()
        r.rt, r.t, r.q = 3, 3, "GT"
        for rd, infos in hits:
            h = (r.a.revcomp_matches if rev else r.a.forward_matches).add()
            h.r = rd
            for g, c, l in infos:
                i = h.s.add()
                i.g, i.c, i.l = g, c, l
        spec = "|".join(rd + ":" + ";".join(f"{g}/{c}/{l}" for g, c, l in infos) for rd, infos in hits)
        cmds.append(f"all\t3\tGT\t{rev}\t{spec}")
        want.append(r.SerializeToString().hex())
    # samples
    r = M["Reply"]()
    r.rt, r.t, r.q = 4, 4, "ACGTAC"
    for g, c, l in [("x", 1, 2), ("yy", 200, 33)]:
        i = r.s.s.add()
        i.g, i.c, i.l = g, c, l
    cmds.append("samples\t4\tACGTAC\tx/1/2;yy/200/33")
    want.append(r.SerializeToString().hex())
    assert _run(exe, cmds) == want


def test_request_round_trip():
    M = proto_schema.build()
    exe = _wire_exe()
    cmds, want = [], []
    for kw in [dict(t=1, rt=1, q="ACGT"), dict(t=2, rt=3, q="ACGTN" * 20, k=31, s=2),
               dict(t=4, rt=4, q="A", p=123456, a="T", isalt=1), dict(t=3, rt=2, q="GATTACA", k=-5)]:
        r = M["Request"](**kw)
        b = r.SerializeToString()
        cmds.append("req\t" + b.hex())
        f = lambda n: str(kw[n]) if n in kw else "-"
        want.append(f"ok {kw['t']} {kw['rt']} {kw['q']} {f('k')} {f('s')} {f('p')} {f('a')} {f('isalt')} {b.hex()}")
    cmds += ["req\t0801", "req\tff", "req\t08011001"]  # required fields missing / truncated varint
    want += ["bad", "bad", "bad"]
    assert _run(exe, cmds) == want


@pytest.mark.gpu
def test_count_batcher_sends_the_reference_replies(tmp_path):
    import oracle
    M = proto_schema.build()
    lib = rs_build.build()
    exe = os.path.join(CPP, "service_test")
    subprocess.run(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(CPP, "service_test.cpp"), "-o", exe, "-L", os.path.dirname(lib), "-lrs_b200",
                    "-Wl,-rpath," + os.path.dirname(lib)], check=True)
    bwt = os.path.join(common.GOLDEN, "clean.bwt")
    z = np.load(os.path.join(common.GOLDEN, "clean.npz"))
    P = oracle.PortIndex(bwt)
    qs = [bytes(k).decode() for k in z["kmers"][:300]] + [str(s) for s in z["var_queries"]] + ["ACGTNACGT", "ACGT", "G"]
    comp = str.maketrans("ACGT", "TGCA")
    reqs, want, skipped = [], [], 0
    for i, q in enumerate(qs):
        t, rt = (1, 1) if i % 3 else (2, 1)          # CountReads, or ExactMatch with output=count
        if i % 50 == 49:
            t, rt = 2, 2                              # ExactMatch/Reads: not the batcher's
            skipped += 1
        reqs.append(M["Request"](t=t, rt=rt, q=q).SerializeToString().hex())
        if (t, rt) == (2, 2):
            continue
        for rev in (False, True):
            s = q[::-1].translate(comp) if rev else q
            if any(c not in "ACGT" for c in s):
                cnt = 0
            else:
                lo, up = P.find_interval(s)
                cnt = up - lo + 1 if up >= lo else 0
            want.append(proto_schema.count_reply(M, t, q, rev, cnt).hex())
    reqs.append("ff")  # garbage on the socket
    skipped += 1
    f = tmp_path / "req.hex"
    f.write_text("\n".join(reqs) + "\n")
    out = subprocess.run([exe, bwt, str(f)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[:skipped] == ["skip"] * skipped
    assert out[skipped] == f"pending {len(qs) - (skipped - 1)}"
    assert out[-1] == f"sent {len(want)}"
    assert out[skipped + 1:-1] == want
    assert sum(1 for w in want if not w.endswith("0800")) > 200  # most counts are non-zero


@pytest.mark.gpu
def test_query_batcher_replies_equal_query_task(tmp_path):
    """rsb200::QueryBatcher against QueryTask::run (service.cpp:1259-1362) restated with google.protobuf over the
    oracle: ExactMatch with output = reads / all / samples for k-mer queries, both strands, rows in find_reads'
    order (an interval of more than 4,096 rows included), sample names through the hash file."""
    import oracle
    M = proto_schema.build()
    lib = rs_build.build()
    exe = os.path.join(CPP, "query_service_test")
    subprocess.run(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(CPP, "query_service_test.cpp"), "-o", exe, "-L", os.path.dirname(lib), "-lrs_b200",
                    "-Wl,-rpath," + os.path.dirname(lib)], check=True)
    bwt = os.path.join(common.GOLDEN, "clean.bwt")
    z = np.load(os.path.join(common.GOLDEN, "clean.npz"))
    P = oracle.PortIndex(bwt)
    gold = common.make_case("clean", *common.Datasets.SPECS["clean"][:3], min_len=73, keep_reads=True)
    sos, nsamples = 1, 40
    rng = np.random.default_rng(4)
    per_read = rng.integers(1, 4, gold.num_strings)
    off = np.zeros(gold.num_strings + 1, dtype=np.uint64)
    off[1:] = np.cumsum(per_read)
    who = rng.integers(0, nsamples, int(off[-1]))
    codes = np.frombuffer(b"".join(oracle.sample_code(int(s), sos) for s in who), dtype=np.uint8)
    etl = str(tmp_path / "merged.rlosorted.reads")
    common.write_etl_file(etl, gold.reads, off, codes, sos, False)
    names = {oracle.sample_code(s, sos): f"SAMPLE{s:03d}" for s in range(nsamples)}
    hashfile = tmp_path / "hash.txt"
    hashfile.write_bytes(b"".join(n.encode() + b"\t" + c + b"\n" for c, n in names.items()))

    # (queries of min_read_length bases and more: tests/test_service_shell.py)
    qs = [bytes(k).decode() for k in z["kmers"][:60]] + ["ACGTNACGT", "ACG", "GATTACA"]
    comp = str.maketrans("ACGT", "TGCA")
    reqs, want, skipped = [], [], 0
    for i, q in enumerate(qs):
        rt = (2, 3, 4)[i % 3]
        t = 2
        if i % 20 == 19:
            t, rt = 1, 1                             # CountReads: not this batcher's
        reqs.append(M["Request"](t=t, rt=rt, q=q).SerializeToString().hex())
        if t != 2:
            skipped += 1
            continue
        for rev in (False, True):
            s = q[::-1].translate(comp) if rev else q
            lo, up = P.find_interval(s)
            rows = []
            if up >= lo:
                start = lo
                while up - start > 4096:
                    start += 2048
                rows = list(range(start, up + 1)) + list(range(lo, start))
            r = M["Reply"]()
            r.rt, r.t, r.q = 2, rt, q
            if rt == 2:
                r.r.SetInParent()
                for row in rows:
                    (r.r.revcomp_matches if rev else r.r.forward_matches).add().r = P.extract_read(row)
            else:
                r.a.SetInParent()
                for row in rows:
                    one = (r.a.revcomp_matches if rev else r.a.forward_matches).add()
                    one.r = P.extract_read(row)
                    ri = P.extract_postfix(row)[1]
                    for c in range(int(off[ri]), int(off[ri + 1])):
                        info = one.s.add()
                        info.g, info.c, info.l = names[oracle.sample_code(int(who[c]), sos)], 0, 0
            want.append(r.SerializeToString().hex())
    f = tmp_path / "req.hex"
    f.write_text("\n".join(reqs) + "\n")
    out = subprocess.run([exe, bwt, str(f), etl, str(hashfile), str(sos)], check=True, capture_output=True,
                         text=True).stdout.splitlines()
    assert out[:skipped] == ["skip"] * skipped
    assert out[skipped] == f"pending {len(qs) - skipped}"
    assert out[-1] == f"sent {len(want)}"
    got = out[skipped + 1:-1]
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g == w
    assert max(len(w) for w in want) > 4096 * 73 * 2   # the 3-mer's reply holds more than 4,096 reads
