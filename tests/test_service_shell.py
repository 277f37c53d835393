"""The worker shell around the batched calls (include/rs_config.hpp, include/rs_service.hpp) and the oracles that were
unpinned after round 1, now anchored on the reference's own code:

* the .cfg parser against libconfig 1.5 as vendored by the reference (oracle/_ref/libref_config.so; golden
  tests/golden/config_golden.json where that library is absent);
* the sample-code arithmetic against src/util/create_samples_hash.cpp (golden tests/golden/sample_codes.json);
* the /gt restatement oracle/gt.py against find_gt_reads compiled from the reference's service.cpp (golden
  tests/golden/gt_golden.json, and live when oracle/_ref/libref_service.so is present);
* on a GPU: the whole receive loop (rsb200::Worker) on a configs[0]-shaped request script -- demo/test.in's query
  lengths, count / reads / all / samples, both strands, middle and long queries, metadata in the sample store --
  against the reference's handlers restated over the oracle with google.protobuf.
"""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import oracle
from oracle import gt as ogt
from readserver_b200 import build as rs_build
from tests import common, proto_schema

ROOT = common.ROOT
CPP = os.path.join(ROOT, "tests", "cpp")
GOLD = common.GOLDEN


def _exe(name, link=False):
    exe = os.path.join(CPP, name)
    cmd = ["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(CPP, name + ".cpp"), "-o", exe]
    if link:
        lib = rs_build.build()
        cmd += ["-L", os.path.dirname(lib), "-lrs_b200", "-Wl,-rpath," + os.path.dirname(lib)]
    subprocess.run(cmd, check=True)
    return exe


def _parse_with_ours(path, keys):
    out = subprocess.run([_exe("config_test"), "parse", path, *keys], capture_output=True, text=True)
    if out.returncode != 0:
        return None
    got = {}
    for line in out.stdout.splitlines():
        f = line.split("\t")
        if f[1] == "-":
            got[f[0]] = None
        elif f[1] == "S":
            got[f[0]] = (f[2] if len(f) > 2 else "").replace("\x01", "\t")
        elif f[1] == "A":
            got[f[0]] = f[2].split("|") if len(f) > 2 and f[2] else []
        else:
            got[f[0]] = "<not a string>"
    return got


def test_config_parser_equals_libconfig(tmp_path):
    from tests.golden.make_golden2 import CFG_ARRAYS, CFG_STRINGS, CFG_TEXT
    want = json.load(open(os.path.join(GOLD, "config_golden.json")))
    path = os.path.join(GOLD, "worker.cfg")
    assert open(path).read() == CFG_TEXT
    got = _parse_with_ours(path, CFG_STRINGS + CFG_ARRAYS)
    assert got == want
    if os.path.exists(oracle.REF_CONFIG_SO):
        assert oracle.ref_config(path, CFG_STRINGS, CFG_ARRAYS) == want
        # more syntax, checked live: ':' assignment, a list in ( ), settings that are not strings, no final newline
        extra = tmp_path / "x.cfg"
        extra.write_text('a : "1"\nb = ( "x", "y" );\nn = 12;\nf = 1.5; t = true\ng = { inner = "v"; };\nz = "end"')
        keys = ["a", "z", "n", "f", "t", "g"]
        ref = oracle.ref_config(str(extra), ["a", "z"], ["b"])
        ours = _parse_with_ours(str(extra), ["a", "z", "b", "n", "f", "t", "g"])
        assert ours["a"] == ref["a"] == "1" and ours["z"] == ref["z"] == "end" and ours["b"] == ref["b"] == ["x", "y"]
        assert all(ours[k] == "<not a string>" for k in ("n", "f", "t", "g")), keys
        tmpl = "/root/reference/demo/TEMPLATE.service.cfg"
        if os.path.exists(tmpl):   # the reference's own template, where the reference tree is present
            strings = ["prefix", "hashfile", "pull", "push", "push_count", "suffix", "size_of_sample", "has_other_meta_data",
                       "max_read_length", "min_read_length", "rocksdb_path", "rocksdb_ext"]
            assert _parse_with_ours(tmpl, strings + ["rocksdb"]) == oracle.ref_config(tmpl, strings, ["rocksdb"])
    lenient = tmp_path / "lenient.cfg"            # what libconfig lets pass, passes here too, with the same value
    lenient.write_text('a = "\\q" /* never closed')
    assert _parse_with_ours(str(lenient), ["a"]) == {"a": "\\q"}
    if os.path.exists(oracle.REF_CONFIG_SO):
        assert oracle.ref_config(str(lenient), ["a"], []) == {"a": "\\q"}
    for bad in ['a = "unterminated', "a = [ \"x\", ", "= \"v\"", 'a = "1"; a = "2";']:
        p = tmp_path / "bad.cfg"
        p.write_text(bad)
        assert _parse_with_ours(str(p), ["a"]) is None, bad
        if os.path.exists(oracle.REF_CONFIG_SO):
            with pytest.raises(ValueError):
                oracle.ref_config(str(p), ["a"], [])
    out = subprocess.run([_exe("config_test"), "worker", path], capture_output=True, text=True, check=True).stdout.split()
    assert out[:4] == ["1", "0", "100", "73"] and out[4] == "/data/bwt/final" and out[5] == "[]" and out[-1] == "4"


def test_batch_window_is_size_or_time():
    out = subprocess.run([_exe("config_test"), "window"], capture_output=True, text=True, check=True).stdout.strip()
    assert out == "01000 0 600 1 0 1 0 01"


def test_sample_codes_equal_create_samples_hash():
    gold = json.load(open(os.path.join(GOLD, "sample_codes.json")))
    for n, g in gold.items():
        n, w = int(n), g["width"]
        mine = b"".join(oracle.sample_code(i, w) for i in range(n))
        assert hashlib.sha256(mine).hexdigest() == g["sha256"], n
        assert [oracle.sample_code(i, w).hex() for i in range(4)] == g["first"]
        if g["all_hex"]:
            assert mine.hex() == g["all_hex"]
    # width rule of the tool: bytes needed for n keys in base 223 (create_samples_hash.cpp:31-39)
    assert [gold[k]["width"] for k in ("5", "222", "223", "1000", "60000")] == [1, 1, 2, 2, 3]
    if os.path.exists(oracle.REF_CREATE_SAMPLES_HASH):
        names = [f"n{i:04d}" for i in range(300)]
        live = oracle.ref_create_samples_hash(names)
        assert [live[x] for x in sorted(names)] == [oracle.sample_code(i, 2) for i in range(300)]
    # decode of a stored value, with and without the count / lane bytes (service.cpp:829-843)
    v = oracle.sample_code(0, 2) + bytes([33 + 5, 33 + 2]) + oracle.sample_code(299, 2) + bytes([33 + 90, 33 + 7])
    assert oracle.decode_sample_value(v, 2, True) == [(oracle.sample_code(0, 2), 5, 2), (oracle.sample_code(299, 2), 90, 7)]


def _digest(strings):
    return hashlib.sha1("\n".join(strings).encode()).hexdigest()


def test_gt_restatement_equals_the_references_find_gt_reads(datasets):
    """oracle/gt.py, which the GPU /gt seed stage is checked against, returns for every tile exactly the reads the
    reference's own find_gt_reads returns: against vectors generated from the reference (always) and against the
    compiled function itself (where oracle/_ref holds it)."""
    gold = json.load(open(os.path.join(GOLD, "gt_golden.json")))
    case, P = datasets.case("clean"), datasets.port("clean")
    live = oracle.RefService(os.path.join(GOLD, "clean.bwt")) if oracle.have_ref_service() else None
    seen = {"reads": 0, "throw": 0, "hang": 0, "tiles": 0}
    for g in gold["windows"]:
        kmers, skip, max_iv, w, pos = g["kmers"], g["skip"], g["max_interval_size"], g["w"], g["pos"]
        if live:
            live.set_limits(max_iv)
        for i, want in enumerate(g["tiles"]):
            seen["tiles"] += 1
            try:
                mine = sorted(case.reads[r] for r in ogt.find_gt_reads(P, w, i, kmers, skip, pos, max_iv))
            except ogt.GtUndefined as e:
                assert want == e.kind, (w, pos, i)
                seen[e.kind] += 1
                if live and e.kind == "throw":
                    assert live.find_gt_reads(w, i, kmers, skip, pos) is None
                continue
            assert want == [len(mine), _digest(mine)], (kmers, skip, max_iv, w, pos, i)
            seen["reads"] += len(mine)
            if live and i % 3 == 0:
                assert live.find_gt_reads(w, i, kmers, skip, pos) == mine
    assert seen["tiles"] > 5000 and seen["reads"] > 20000 and seen["throw"] and seen["hang"]
    if live:   # the little helpers next to it
        assert live.get_tiles("ACGTACGTAC", 4, 1) == sorted({"ACGT", "GTAC"}) and live.get_tiles("ACG", 4) == []
        assert live.rev_comp("AACGTN") == "NACGTT" and live.is_suffix_of("GT", "ACGT") and not live.is_suffix_of("GA", "ACGT")


# ------------------------------------------------------------------------------------------------
# the receive loop on a configs[0]-shaped script
# ------------------------------------------------------------------------------------------------
COMP = str.maketrans("ACGT", "TGCA")


def _find_reads(P, reads, index_of, w, suffix, mn, mx):
    """find_reads (service.cpp:714-797) over the oracle: list of (read string, read index)."""
    sz = len(w)
    acgt = all(c in "ACGT" for c in w)

    def rows_of(lo, up, chunked):
        if up < lo:
            return []
        if not chunked:
            return list(range(lo, up + 1))
        start = lo
        while up - start > 4096:
            start += 2048
        return list(range(start, up + 1)) + list(range(lo, start))

    def tiles(length):
        out, seen = [], set()
        for i in range(0, sz - length + 1):
            t = w[i:i + length]
            if t in seen:
                continue
            seen.add(t)
            if (suffix == "" or t.endswith(suffix)) and t in index_of:
                out.append((t, index_of[t]))
        return out

    def by_rows(rows):
        return [(P.extract_read(r), P.extract_postfix(r)[1]) for r in rows]

    if sz < mn:
        lo, up = P.find_interval(w)
        return by_rows(rows_of(lo, up, True))
    if sz < mx:
        out = tiles(mn) if sz != mn else []
        if acgt:
            lo, up = P.find_interval(w)
            out += by_rows(rows_of(lo, up, False))
        return out
    return tiles(mx) + (tiles(mn) if mn != mx else [])


@pytest.mark.gpu
@pytest.mark.parametrize("meta,suffix", [(False, ""), (True, "A")])
def test_worker_loop_on_a_demo_shaped_script(meta, suffix, tmp_path):
    M = proto_schema.build()
    exe = _exe("worker_test", link=True)
    bwt = os.path.join(GOLD, "clean.bwt")
    P = oracle.PortIndex(bwt)
    gold = common.make_case("clean", *common.Datasets.SPECS["clean"][:3], min_len=73, keep_reads=True)
    index_of = {r: i for i, r in enumerate(gold.reads)}
    sos, nsamples = 2, 300
    rng = np.random.default_rng(8)
    per_read = rng.integers(1, 4, gold.num_strings)
    off = np.zeros(gold.num_strings + 1, dtype=np.uint64)
    off[1:] = np.cumsum(per_read)
    who = rng.integers(0, nsamples, int(off[-1]))
    codes = np.frombuffer(b"".join(oracle.sample_code(int(s), sos) for s in who), dtype=np.uint8)
    etl = str(tmp_path / "merged.rlosorted.reads")
    common.write_etl_file(etl, gold.reads, off, codes, sos, meta, n_reads_every=97)
    # the metadata bytes write_etl_file drew, re-read from the file it wrote
    values = [l for l in open(etl, "rb").read().split(b"\n")[1::2]]
    kept = [v for r, v in zip(open(etl, "rb").read().split(b"\n")[0::2], values) if b"N" not in r and r]
    assert len(kept) == gold.num_strings
    names = {oracle.sample_code(s, sos): f"SAMPLE{s:03d}" for s in range(nsamples)}
    hashfile = tmp_path / "hash.txt"
    hashfile.write_bytes(b"".join(n.encode() + b"\t" + c + b"\n" for c, n in names.items()))
    cfg = tmp_path / "service.cfg"
    cfg.write_text(f'// worker of the stand-in shard\nprefix = "{bwt[:-4]}";\nhashfile = "{hashfile}";\n'
                   f'pull = "tcp://127.0.0.1:1"; push = "tcp://127.0.0.1:2"; push_count = "tcp://127.0.0.1:3";\n'
                   f'suffix = "{suffix}";\nsize_of_sample = "{sos}";\nhas_other_meta_data = "{int(meta)}";\n'
                   f'max_read_length = "100";\nmin_read_length = "73";\nrocksdb_path = "/nowhere/"; rocksdb_ext = ".rocksdb";\n'
                   f'rocksdb = [ "AAA", "AAC" ];\n')

    full = [r for r in gold.reads if len(r) == 100]
    r73 = [r for r in gold.reads if len(r) == 73]
    g = bytes(__import__("readserver_b200.datagen", fromlist=["x"]).codes_to_ascii(gold.genome)).decode()
    # demo/test.in: 9 queries of 37, 41, 46, 41 and 5 x 100 bases; here cut from reads / the genome so that they hit
    qs = [full[3][:37], full[5][10:51], g[1000:1046], full[9][59:100], full[11], full[12], g[5000:5100], full[13], full[14]]
    qs += [r73[0], r73[1] + "ACG", g[700:780], g[2000:2120], full[20][5:95], "ACGTNACGT", "ACG", full[2][:50] + "N" + full[2][51:90]]
    reqs, want = [], []
    for i, q in enumerate(qs):
        kinds = [(1, 1), (2, 1), (2, 2), (2, 3)] if i < 12 else [(2, 2 + i % 2)]
        for t, rt in kinds:
            reqs.append(M["Request"](t=t, rt=rt, q=q).SerializeToString().hex())
            for rev in (False, True):
                s = q[::-1].translate(COMP) if rev else q
                if rt == 1:
                    if any(c not in "ACGT" for c in s):
                        cnt = 0
                    else:
                        lo, up = P.find_interval(s)
                        cnt = up - lo + 1 if up >= lo else 0
                    want.append(("count" if t == 1 else "push", proto_schema.count_reply(M, t, q, rev, cnt).hex()))
                    continue
                hits = _find_reads(P, gold.reads, index_of, s, suffix, 73, 100)
                r = M["Reply"]()
                r.rt, r.t, r.q = 2, rt, q
                if rt == 2:
                    r.r.SetInParent()
                    for rd, _ in hits:
                        (r.r.revcomp_matches if rev else r.r.forward_matches).add().r = rd
                else:
                    r.a.SetInParent()
                    for rd, ri in hits:
                        one = (r.a.revcomp_matches if rev else r.a.forward_matches).add()
                        one.r = rd
                        val = kept[ri]
                        stride = sos + (2 if meta else 0)
                        for c in range(0, len(val), stride):
                            info = one.s.add()
                            info.g = names[val[c:c + sos]]
                            info.c, info.l = (val[c + sos] - 33, val[c + sos + 1] - 33) if meta else (0, 0)
                want.append(("push", r.SerializeToString().hex()))
    # k-mer match (KmerTask, service.cpp:851-990): the reads of all tiles, each once; count = how many
    for q, kk, ss in ((qs[4], 31, 0), (qs[2], 31, 0), (qs[11], 40, 7), (full[2][:50] + "N" + full[2][51:90], 20, 3), (qs[15], 9, 0)):
        for rt in (1, 2, 3):
            reqs.append(M["Request"](t=3, rt=rt, q=q, k=kk, s=ss).SerializeToString().hex())
            for rev in (False, True):
                s = q[::-1].translate(COMP) if rev else q
                hits, seen, tiles_seen = [], set(), set()
                for at in range(0, len(s) - kk + 1, ss + 1):
                    tile = s[at:at + kk]
                    if tile in tiles_seen or any(c not in "ACGT" for c in tile):
                        continue
                    tiles_seen.add(tile)
                    for rd, ri in _find_reads(P, gold.reads, index_of, tile, suffix, 73, 100):
                        if rd not in seen:
                            seen.add(rd)
                            hits.append((rd, ri))
                r = M["Reply"]()
                r.rt, r.t, r.q = 3, rt, q
                if rt == 1:
                    (r.c.revcomp_matches if rev else r.c.forward_matches).c = len(hits)
                elif rt == 2:
                    r.r.SetInParent()
                    for rd, _ in hits:
                        (r.r.revcomp_matches if rev else r.r.forward_matches).add().r = rd
                else:
                    r.a.SetInParent()
                    for rd, ri in hits:
                        one = (r.a.revcomp_matches if rev else r.a.forward_matches).add()
                        one.r = rd
                        val = kept[ri]
                        stride = sos + (2 if meta else 0)
                        for c in range(0, len(val), stride):
                            info = one.s.add()
                            info.g = names[val[c:c + sos]]
                            info.c, info.l = (val[c + sos] - 33, val[c + sos + 1] - 33) if meta else (0, 0)
                want.append(("push", r.SerializeToString().hex()))
    kmer_hits = sum(len(M["Reply"].FromString(bytes.fromhex(h)).r.forward_matches) for s_, h in want
                    if M["Reply"].FromString(bytes.fromhex(h)).rt == 3)
    assert kmer_hits > 10
    # site match (GtTask, service.cpp:993-1212): window cut around p, seeds by find_gt_reads (oracle/gt.py, pinned on the
    # reference's own), the aligner a stand-in predicate (the same in tests/cpp/worker_test.cpp), count / reads / all
    from oracle import gt as ogt
    site_reqs = [(g[3000:3260], 130, "A", 0, 31, 7), (g[8000:8150], 40, "ACG", 1, 31, 7), (g[500:700], 190, "T", 0, 20, 3),
                 (g[100:180], 300, "A", 0, 31, 7), ("ACGT", 2, "A", 0, 31, 7), (full[30] + g[9000:9100], 100, "G", 1, 31, 0)]
    site_hits = 0
    for q, p_, a, isalt, kk, ss in site_reqs:
        for rt in (1, 2, 3):
            reqs.append(M["Request"](t=4, rt=rt, q=q, k=kk, s=ss, p=p_, a=a, isalt=isalt).SerializeToString().hex())
            for rev in (False, True):
                w, pos, alt = q, p_, a
                if rev:
                    w, pos, alt = q[::-1].translate(COMP), len(q) - p_ + 1, a[::-1].translate(COMP)
                bad = pos < 0
                if not bad and pos <= 100:
                    w = w[:100 + pos - 1]
                elif not bad and pos - 100 > len(w):
                    bad = True
                elif not bad:
                    w, pos = w[pos - 100: pos - 100 + 199], 100
                bad = bad or pos > len(w) or len(w) < kk
                r = M["Reply"]()
                r.rt, r.t, r.q = 4, rt, q
                if not bad:
                    idx, undefined = ogt.gt_task_reads(P, w, pos, kk, ss, 10000)
                    bad = bool(undefined)
                if not bad:
                    keep = [ri for ri in idx
                            if (len(gold.reads[ri]) + ord(gold.reads[ri][len(gold.reads[ri]) // 2]) + len(w) + pos + len(alt) + isalt) % 3]
                    site_hits += len(keep)
                    if rt == 1:
                        (r.c.revcomp_matches if rev else r.c.forward_matches).c = len(keep)
                    elif rt == 2:
                        r.r.SetInParent()
                        for ri in keep:
                            (r.r.revcomp_matches if rev else r.r.forward_matches).add().r = gold.reads[ri]
                    else:
                        r.a.SetInParent()
                        for ri in keep:
                            one = (r.a.revcomp_matches if rev else r.a.forward_matches).add()
                            one.r = gold.reads[ri]
                            val = kept[ri]
                            stride = sos + (2 if meta else 0)
                            for c in range(0, len(val), stride):
                                info = one.s.add()
                                info.g = names[val[c:c + sos]]
                                info.c, info.l = (val[c + sos] - 33, val[c + sos + 1] - 33) if meta else (0, 0)
                want.append(("site", r.SerializeToString().hex()))
    assert site_hits > 30
    # nobody's: garbage, a k-mer request without k, a site request without k
    others = [M["Request"](t=3, rt=2, q=qs[4], s=0).SerializeToString().hex(),
              M["Request"](t=4, rt=3, q=qs[5], s=7, p=40).SerializeToString().hex(), "ff"]
    reqs[5:5] = others[:1]
    reqs += others[1:]
    f = tmp_path / "req.hex"
    f.write_text("\n".join(reqs) + "\n")
    sai = tmp_path / "shard.sai"
    np.asarray(gold.sai, dtype=np.uint64).tofile(sai)
    out = subprocess.run([exe, str(cfg), str(f), etl, str(sai)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[-1] == f"sent {len(want)}"
    got_other = [l.split(" ", 1)[1] for l in out if l.startswith("other ")]
    assert got_other == others
    # replies of one sink and one kind arrive in request order; batches of the two kinds interleave by their windows
    got = [tuple(l.split(" ", 1)) for l in out[:-1] if not l.startswith("other ")]
    # (the count batcher answers ExactMatch counts, the query batcher everything about reads and all of KmerMatch)
    def is_site(h):
        return M["Reply"].FromString(bytes.fromhex(h)).rt == 4
    assert [h for s_, h in got if s_ == "push" and is_site(h)] == [h for s_, h in want if s_ == "site"]
    got = [(s_, h) for s_, h in got if not is_site(h)]
    for sink, is_count in (("count", True), ("push", True), ("push", False)):
        def kind(h):
            m = M["Reply"].FromString(bytes.fromhex(h))
            return m.HasField("c") and m.rt != 3
        g_ = [h for s_, h in got if s_ == sink and kind(h) == is_count]
        w_ = [h for s_, h in want if s_ == sink and kind(h) == is_count]
        assert g_ == w_, (sink, is_count)
    nonempty = sum(1 for s_, h in want if s_ == "push" and len(h) > 400)
    assert nonempty > 20
