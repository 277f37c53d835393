"""GPU parity of this round's additions, through the C ABI against the CPU oracle:
the device-side re-encoder (K4), the persisted index, and the shard-set pipeline (count + locate + sample sets over
the reference's 64 read-suffix shards; front end fan-in server.cpp:184-197, 404-410)."""
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
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from readserver_b200 import engine as e
    return e


# ------------------------------------------------------------------------------------------------
# K4 on the device == K4 on the host, array by array
# ------------------------------------------------------------------------------------------------
def _arrays_equal(a, b):
    for what in range(0, 8):
        assert np.array_equal(a.export(what), b.export(what)), what
    ia, ib = a.info, b.info
    for f in ("bwlen", "num_strings", "num_runs", "num_blocks", "ref_depth", "ref_clamp_c", "ref_vlast", "q2_safe"):
        assert getattr(ia, f) == getattr(ib, f), f
    assert list(ia.pc) == list(ib.pc) and list(ia.total) == list(ib.total)


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_device_encoder_equals_host_encoder(name, datasets, engine):
    import torch
    case = datasets.case(name)
    kw = dict(num_strings=case.num_strings, num_symbols=case.num_symbols)
    with engine.GpuBWT(runs=case.runs, **kw) as D, engine.GpuBWT(runs=case.runs, host_encoder=True, **kw) as H, \
            engine.GpuBWT(runs=torch.from_numpy(case.runs).cuda(), **kw) as DD:
        _arrays_equal(D, H)
        _arrays_equal(DD, H)
        # getChar's own clamp is not in rs_index_info: compare its effect at the end of the BWT
        idx = np.arange(max(0, case.num_symbols - 70000), case.num_symbols, dtype=np.uint64)
        assert np.array_equal(D.get_char(idx), H.get_char(idx))
        q = case.queries(2000, 31, seed=77)
        assert all(np.array_equal(x, y) for x, y in zip(D.find_interval(q, 31), H.find_interval(q, 31)))


@pytest.mark.parametrize("shape", ["ones", "single", "multiple36", "rand_depth3", "rand_small"])
def test_device_encoder_on_degenerate_run_shapes(shape, engine):
    rng = np.random.default_rng(12)
    if shape == "ones":          # every run has length 1: more than three block boundaries per 512-symbol slot
        runs = ((rng.integers(0, 5, 20000) << 5) | 1).astype(np.uint8)
    elif shape == "single":
        runs = np.array([(2 << 5) | 17], dtype=np.uint8)
    elif shape == "multiple36":
        runs = ((rng.integers(1, 5, 36 * 50) << 5) | rng.integers(1, 32, 36 * 50)).astype(np.uint8)
    elif shape == "rand_depth3":
        runs, _ = random_runs(5, 1_200_000)
    else:
        runs, _ = random_runs(6, 70_000)
    n = int((runs & 31).sum())
    ns = int((runs[(runs >> 5) == 0] & 31).sum())
    with engine.GpuBWT(runs=runs, num_strings=ns, num_symbols=n) as D, \
            engine.GpuBWT(runs=runs, num_strings=ns, num_symbols=n, host_encoder=True) as H:
        _arrays_equal(D, H)
        P = oracle.PortIndex(runs=runs, num_strings=ns, num_symbols=n)
        idx = common.positions(n, limit=100_000)
        for c in b"$ACGT":
            s = np.full(idx.size, c, dtype=np.uint8)
            assert np.array_equal(D.get_occ(s, idx), P.get_occ_batch(s, idx))
        assert np.array_equal(D.get_char(idx), P.get_char_batch(idx))


def test_device_encoder_rejects_what_the_host_encoder_rejects(engine):
    bad = np.array([(5 << 5) | 3, (1 << 5) | 2], dtype=np.uint8)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(runs=bad, num_strings=0, num_symbols=5)
    assert e.value.status == 4
    zero = np.array([(1 << 5) | 0], dtype=np.uint8)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(runs=zero, num_strings=0, num_symbols=0)
    assert e.value.status == 4
    ok = np.array([(1 << 5) | 3], dtype=np.uint8)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(runs=ok, num_strings=0, num_symbols=4)   # symbol count disagrees with the runs
    assert e.value.status == 5


@pytest.mark.parametrize("name", ["small", "deep"])
def test_saved_index_reopens_identically(name, datasets, engine, tmp_path):
    case, P = datasets.case(name), datasets.port(name)
    path = str(tmp_path / "shard.rsb2")
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        B.set_sai(case.sai)
        B.save(path)
        with engine.GpuBWT(encoded_path=path) as R:
            _arrays_equal(B, R)
            q = case.queries(5000, 31, seed=3)
            lo, up = R.find_interval(q, 31)
            lo2, up2 = P.find_interval_batch(q, 31, nthreads=4)
            assert np.array_equal(lo, lo2) and np.array_equal(up, up2)
            hit = np.nonzero(up >= lo)[0][:300]
            rows = np.concatenate([np.arange(lo[i], up[i] + 1, dtype=np.uint64) for i in hit])
            a = R.extract_prefix(rows, max_steps=128, want_read_index=True)
            b = B.extract_prefix(rows, max_steps=128, want_read_index=True)
            assert all(np.array_equal(x, y) for x, y in zip(a, b))
    with open(path, "r+b") as f:          # truncated and foreign files are refused with a status, not a crash
        f.truncate(os.path.getsize(path) // 2)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(encoded_path=path)
    assert e.value.status == 3
    with open(path, "wb") as f:
        f.write(b"not an index" * 100)
    with pytest.raises(engine.RsError) as e:
        engine.GpuBWT(encoded_path=path)
    assert e.value.status == 2


# ------------------------------------------------------------------------------------------------
# shard set: count + locate + sample sets over the 64 read-suffix shards
# ------------------------------------------------------------------------------------------------
class PopCase:
    """A small population (configs[2] shape), its 64 shards on the GPU and their CPU oracles."""

    def __init__(self, engine, genome_len=400_000, n_samples=9, coverage=20, read_len=100, device_runs=True, device=0):
        import torch
        self.pop = dg.make_population(genome_len, n_samples, seed=4, device=f"cuda:{device}")
        keys, self.hist = dg.population_read_keys(self.pop, coverage, read_len, seed=41)
        self.read_len = read_len
        self.gpu, self.port, self.sai, self.samp, self.reads = [], [], [], [], []
        for s in range(64):
            sh = dg.build_population_shard(self.pop, keys, s, read_len)
            b = sh.built
            runs_h = b.runs.cpu().numpy()
            B = engine.GpuBWT(runs=b.runs.contiguous() if device_runs else runs_h, num_strings=b.num_strings,
                              num_symbols=b.num_symbols, device=device)
            B.set_sai(b.sai.contiguous() if device_runs else b.sai.cpu().numpy().astype(np.uint64))
            if device_runs:
                B.set_samples(sh.samp_off.contiguous(), sh.samp_codes.contiguous(), 1)
            else:
                B.set_samples(sh.samp_off.cpu().numpy().astype(np.uint64), sh.samp_codes.cpu().numpy(), 1)
            self.gpu.append(B)
            self.port.append(oracle.PortIndex(runs=runs_h, num_strings=b.num_strings, num_symbols=b.num_symbols))
            self.sai.append(b.sai.cpu().numpy().astype(np.uint64))
            self.samp.append((sh.samp_off.cpu().numpy(), sh.samp_codes.cpu().numpy()))
            self.reads.append(dg.ASCII[sh.reads.cpu().numpy()])
        self.queries = dg.population_queries(self.pop, 3000, 31, seed=5).cpu().numpy()

    def expected(self, q, max_hits, shard_order=None):
        """counts, hit_offsets, hits, carriers (4 words) from the per-shard oracles, in the set's hit order."""
        order = list(range(64)) if shard_order is None else shard_order
        Q = q.shape[0]
        per = []
        counts = np.zeros(Q, dtype=np.uint64)
        for s in order:
            lo, up = self.port[s].find_interval_batch(q, q.shape[1], nthreads=4)
            c = np.where(up >= lo, up - lo + 1, 0).astype(np.uint64)
            counts += c
            per.append((s, lo, np.minimum(c, max_hits)))
        hits = [[] for _ in range(Q)]
        for s, lo, c in per:
            nz = np.nonzero(c)[0]
            if nz.size == 0:
                continue
            rows = np.concatenate([np.arange(lo[i], lo[i] + c[i], dtype=np.uint64) for i in nz])
            lex, _, _, _ = self.port[s].locate_batch(rows, prefix=True, postfix=False, max_steps=self.read_len + 28)
            ns = self.sai[s].size
            ri = np.where(lex < ns, self.sai[s][np.minimum(lex, ns - 1)], U64MAX)
            at = 0
            for i in nz:
                for r in ri[at: at + int(c[i])]:
                    hits[i].append(U64MAX if r == U64MAX else (np.uint64(s) << np.uint64(40)) | r)
                at += int(c[i])
        off = np.zeros(Q + 1, dtype=np.uint64)
        off[1:] = np.cumsum([len(h) for h in hits])
        flat = np.array([x for h in hits for x in h], dtype=np.uint64)
        carr = np.zeros((Q, 4), dtype=np.uint32)
        for i, h in enumerate(hits):
            for x in h:
                if x == U64MAX:
                    continue
                s, r = int(x) >> 40, int(x) & ((1 << 40) - 1)
                so, sc = self.samp[s]
                for code in sc[so[r]: so[r + 1]]:
                    smp = int(code) - 34
                    carr[i, smp >> 5] |= np.uint32(1 << (smp & 31))
        return counts, off, flat, carr

    def close(self):
        for b in self.gpu:
            b.close()


@pytest.fixture(scope="module")
def popcase(engine):
    c = PopCase(engine)
    yield c
    c.close()


@pytest.mark.parametrize("locate_every,mode", [(1, "query"), (1, "shard"), (8, None), (0, None)])
def test_shard_set_count_locate_carriers_match_per_shard_oracles(locate_every, mode, popcase, engine, monkeypatch):
    if mode:
        monkeypatch.setenv("RS_B200_GATHER_MODE", mode)
    q = popcase.queries
    with engine.ShardSet() as S:
        for s, b in enumerate(popcase.gpu):
            S.add(b, s)
        info = S.seal(popcase.read_len + 28, locate_every)
        assert info.num_shards == 64 and info.num_samples == 9 and info.carrier_words == 4
        for max_hits in (1 << 20, 3):
            want = popcase.expected(q, max_hits)
            got = S.query(q, 31, max_hits=max_hits, carriers=True)
            assert np.array_equal(got["counts"], want[0])
            assert np.array_equal(got["hit_offsets"], want[1])
            assert np.array_equal(got["hits"], want[2])
            assert np.array_equal(got["carriers"], want[3])
        # shards of this size have depth-2 trees: the reference loses most matches to its tail behaviour (SURVEY.md Q1:
        # initInterval's getOcc(b, n-1) is short), and so must we; the exact-rank test below sees all of them
        assert int(want[0].sum()) > q.shape[0]
        # the 5-byte hit list of rs_pop_query_hits40 widens to the same list (2^40 - 1 <-> 2^64 - 1)
        got40 = S.query(q, 31, max_hits=3, carriers=True, hits40=True)
        assert all(np.array_equal(got[k], got40[k]) for k in got)
        # packed queries give the same answers; count-only call
        got2 = S.query(engine.pack_kmers(q), 31, packed=True, max_hits=3, carriers=True)
        assert all(np.array_equal(got[k], got2[k]) for k in got)
        assert np.array_equal(S.query(q, 31, locate=False)["counts"], want[0])
        # the reverse-complement task of every request (service.cpp:289-291): ASCII and packed keys
        comp = np.zeros(256, dtype=np.uint8)
        for a, b in zip(b"ACGT", b"TGCA"):
            comp[a] = b
        qr = np.ascontiguousarray(comp[q[:, ::-1]])
        fwd_of_rc = S.query(qr, 31, max_hits=3, carriers=True)
        rc1 = S.query(q, 31, max_hits=3, carriers=True, strand=1)
        rc2 = S.query(engine.pack_kmers(q), 31, max_hits=3, carriers=True, strand=1, packed=True)
        for k_ in fwd_of_rc:
            assert np.array_equal(fwd_of_rc[k_], rc1[k_]) and np.array_equal(fwd_of_rc[k_], rc2[k_]), k_


def test_shard_set_with_exact_rank_equals_brute_force_over_the_reads(popcase, engine):
    """Tail emulation off: counts = substring occurrences over every shard's reads; a k-mer's hits = exactly the reads
    holding it (one hit per occurrence); its carrier set = the samples of those reads."""
    q = popcase.queries[:120]
    try:
        for b in popcase.gpu:
            b.set_tail_emulation(False)
        with engine.ShardSet() as S:
            for s, b in enumerate(popcase.gpu):
                S.add(b, s)
            S.seal(popcase.read_len + 28, 1)
            got = S.query(q, 31, carriers=True)
        strs = [[r.tobytes().decode() for r in popcase.reads[s]] for s in range(64)]
        total = 0
        for i in range(q.shape[0]):
            kmer = q[i].tobytes().decode()
            want_hits, want_carr = [], np.zeros(4, dtype=np.uint32)
            for s in range(64):
                so, sc = popcase.samp[s]
                for r, read in enumerate(strs[s]):
                    at = read.find(kmer)
                    while at >= 0:
                        want_hits.append((s << 40) | r)
                        for code in sc[so[r]: so[r + 1]]:
                            want_carr[(int(code) - 34) >> 5] |= np.uint32(1 << ((int(code) - 34) & 31))
                        at = read.find(kmer, at + 1)
            a, b_ = int(got["hit_offsets"][i]), int(got["hit_offsets"][i + 1])
            assert int(got["counts"][i]) == len(want_hits) == b_ - a
            assert sorted(int(x) for x in got["hits"][a:b_]) == sorted(want_hits)
            assert np.array_equal(got["carriers"][i], want_carr)
            total += len(want_hits)
        assert total > 50 * q.shape[0]
    finally:
        for b in popcase.gpu:
            b.set_tail_emulation(True)


def test_shard_set_order_ids_chunks_and_bad_bytes(popcase, engine, monkeypatch):
    """Shards added in another order with other ids; a chunk size that splits the batch unevenly; k-mers holding a
    byte outside ACGT count 0 (service.cpp:299-301); k < 31."""
    monkeypatch.setenv("RS_B200_POP_CHUNK", "1100")
    order = [63 - s for s in range(0, 64, 3)]
    q = popcase.queries[:2500, :24].copy()
    q[::50, 7] = ord("N")
    with engine.ShardSet() as S:
        for s in order:
            S.add(popcase.gpu[s], s)
        S.seal(popcase.read_len + 28, 1)
        assert S.info.chunk_queries == 1100
        want = popcase.expected(q, 1 << 20, shard_order=order)
        assert int(want[0][::50].sum()) == 0
        got = S.query(q, 24, carriers=True)
        assert np.array_equal(got["counts"], want[0]) and np.array_equal(got["hit_offsets"], want[1])
        assert np.array_equal(got["hits"], want[2]) and np.array_equal(got["carriers"], want[3])
        # capacity too small: offsets and total still reported, status CAPACITY
        import ctypes as C
        L = engine.lib()
        counts = np.zeros(q.shape[0], dtype=np.uint64)
        off = np.zeros(q.shape[0] + 1, dtype=np.uint64)
        hits = np.zeros(10, dtype=np.uint64)
        tot = C.c_uint64()
        flat = np.ascontiguousarray(q).reshape(-1)
        rc = L.rs_pop_query(S._h, flat.ctypes.data, 24, q.shape[0], 0, 0, 1 << 20, counts.ctypes.data, off.ctypes.data,
                            hits.ctypes.data, 10, None, C.byref(tot))
        assert rc == 9 and tot.value == want[1][-1] and np.array_equal(off, want[1]) and np.array_equal(hits, want[2][:10])


@pytest.mark.parametrize("nshards,lanes,pass_bytes", [
    (3, None, None), (7, None, None), (9, None, None), (13, None, None), (16, None, None), (17, None, None), (22, None, None),
    (13, 1, None), (13, 4, None), (7, 32, None), (40, 16, None), (5, 2, None),
    (64, None, 1), (22, None, 1), (40, 4, 30000), (64, None, 0),
    (64, None, "shard"), (9, None, "shard"), (1, None, "shard"), (64, None, "shard-unsorted")])
def test_shard_set_of_few_shards(nshards, lanes, pass_bytes, popcase, engine, monkeypatch):
    """A rank of a multi-GPU job holds a few of the 64 shards: the gather kernel then works on several queries per warp
    (1..16 lanes each, picked from the number of shards; forced here as well).  Long k-mers (a few rows per shard: one
    lane per shard) and short ones (intervals of many rows: the whole warp reads one shard's rows)."""
    if lanes is not None:
        monkeypatch.setenv("RS_B200_GATHER_LANES", str(lanes))
    if pass_bytes in ("shard", "shard-unsorted"):   # the gather shard by shard, one lane per query, over the sorted batch
        monkeypatch.setenv("RS_B200_GATHER_MODE", "shard")   # (what large batches get)
        if pass_bytes == "shard-unsorted":
            monkeypatch.setenv("RS_B200_POP_SORT", "0")
    elif pass_bytes is not None:  # the gather in passes of 8..32 shards whose sample-set tables fit this many bytes
        monkeypatch.setenv("RS_B200_POP_PASS_BYTES", str(pass_bytes))
    order = [(5 * s + 2) % 64 for s in range(nshards)]
    with engine.ShardSet() as S:
        for s in order:
            S.add(popcase.gpu[s], s)
        S.seal(popcase.read_len + 28, 1)
        for k, nq, max_hits in ((31, 3001, 1 << 20), (9, 700, 1 << 20), (9, 700, 11)):
            q = np.ascontiguousarray(popcase.queries[:nq, :k])
            want = popcase.expected(q, max_hits, shard_order=order)
            got = S.query(q, k, max_hits=max_hits, carriers=True)
            assert np.array_equal(got["counts"], want[0]) and np.array_equal(got["hit_offsets"], want[1]), (k, max_hits)
            assert np.array_equal(got["hits"], want[2]) and np.array_equal(got["carriers"], want[3]), (k, max_hits)
        assert int(want[0].max()) > 8 * nshards  # the short k-mers do have long intervals


def test_hits40_sub_batches_and_limits(popcase, engine, monkeypatch):
    """rs_pop_query_hits40: stretches of the list packed sub-batch by sub-batch (boundaries that cut a group of eight
    hits), cut-off walks as 2^40 - 1, shard ids that do not fit refused."""
    monkeypatch.setenv("RS_B200_POP_SUB", "1000")
    q = popcase.queries[:5003]
    with engine.ShardSet() as S:
        for s, b in enumerate(popcase.gpu):
            S.add(b, s)
        S.seal(popcase.read_len + 28, 1)
        a, b_ = S.query(q, 31, carriers=True), S.query(q, 31, carriers=True, hits40=True)
        assert all(np.array_equal(a[k], b_[k]) for k in a) and a["hits"].size > 5000
    with engine.ShardSet() as S:          # plain walks bounded at 3 steps: most hits come back as "no read"
        for s, b in enumerate(popcase.gpu[:5]):
            S.add(b, s)
        S.seal(3, 0)
        a, b_ = S.query(q, 31), S.query(q, 31, hits40=True)
        assert np.array_equal(a["hits"], b_["hits"]) and int((a["hits"] == U64MAX).sum()) > 100
    with engine.ShardSet() as S:
        S.add(popcase.gpu[0], 300)        # a shard id the 40-bit form cannot carry
        S.seal(popcase.read_len + 28, 1)
        assert S.query(q, 31)["hits"].size > 0
        with pytest.raises(engine.RsError) as e:
            S.query(q, 31, hits40=True)
        assert e.value.status == 7 or "40 bits" in str(e.value)


@pytest.mark.parametrize("fill", [True, False])
def test_shard_set_under_the_reference_tail_behaviour(fill, datasets, engine, monkeypatch):
    """A single shard whose tail region is NOT empty ('small': walks into it are cut or wander): the reads' walks leave
    those rows of the read-index table open; the seal walks them once (fill) or the pipeline walks them per query
    (RS_B200_NO_RID_FILL), ending where K2 and the oracle end either way."""
    if not fill:
        monkeypatch.setenv("RS_B200_NO_RID_FILL", "1")
        monkeypatch.setenv("RS_B200_GATHER_MODE", "shard")   # (the walk fallback of the shard-major kernel)
    case, P = datasets.case("small"), datasets.port("small")
    q = case.queries(4000, 20, seed=9)
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        B.set_sai(case.sai)
        for every in (1, 4, 0):
            with engine.ShardSet() as S:
                S.add(B, 17)
                S.seal(128, every)
                got = S.query(q, 20, max_hits=64)
                lo, up = P.find_interval_batch(q, 20, nthreads=4)
                c = np.where(up >= lo, up - lo + 1, 0).astype(np.uint64)
                assert np.array_equal(got["counts"], c)
                ch = np.minimum(c, 64)
                rows = np.concatenate([np.arange(lo[i], lo[i] + ch[i], dtype=np.uint64) for i in np.nonzero(ch)[0]])
                lex, _, _, _ = P.locate_batch(rows, prefix=True, postfix=False, max_steps=128)
                ns = case.sai.size
                ri = np.where(lex < ns, case.sai[np.minimum(lex, ns - 1)], U64MAX)
                want = np.where(ri == U64MAX, U64MAX, (np.uint64(17) << np.uint64(40)) | ri)
                assert np.array_equal(got["hits"], want), every


def test_shard_set_argument_errors(popcase, engine):
    with engine.ShardSet() as S:
        with pytest.raises(engine.RsError):
            S.seal(100, 1)                       # empty
        S.add(popcase.gpu[0], 0)
        with pytest.raises(engine.RsError):
            S.add(popcase.gpu[1], 0)             # id twice
        with pytest.raises(engine.RsError):
            S.query(popcase.queries[:4], 31)     # not sealed
        S.seal(128, 1)
        with pytest.raises(engine.RsError):
            S.query(np.zeros((2, 40), dtype=np.uint8) + 65, 40)   # k > 31


@pytest.mark.parametrize("locate_every,world,mode,chunk", [(1, 3, "query", "1024"), (1, 2, "shard", "1024"),
                                                          (1, 3, "shard", None), (0, 2, None, "1024")])
def test_multi_rank_fan_in_through_peer_memory(locate_every, world, mode, chunk):
    """The whole multi-GPU protocol with `world` PROCESSES (tests/pop_worker.py): CUDA IPC mapping of every rank's result
    buffers, counts added into the owner's buffer, hit numbers, slots assigned by the owner, read indices written into
    the owner's list, sample sets OR-ed into the owner's bitmap, flag barriers in between.  The ranks use GPU
    rank % device_count, so on a one-GPU box they share the device (time-sliced); every owner must hold exactly what one
    shard set over all 64 shards answers for its k-mers, the hits in rank order."""
    import socket
    import subprocess
    import sys
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   **({"RS_B200_POP_CHUNK": chunk} if chunk else {}),   # None: one chunk, so the batch is sorted
                   **({"RS_B200_GATHER_MODE": mode} if mode else {}))
        procs.append(subprocess.Popen([sys.executable, os.path.join(common.ROOT, "tests", "pop_worker.py"), str(locate_every)],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for x in procs:
                x.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and "POP_WORKER_OK" in out, f"rank {r}:\n{out[-3000:]}"


# ------------------------------------------------------------------------------------------------
# dense rank blocks: same answers, array checked against the plain BWT
# ------------------------------------------------------------------------------------------------
def _expected_dense(runs: np.ndarray) -> np.ndarray:
    sym = np.repeat(runs >> 5, runs & 31).astype(np.int64)
    n = sym.size
    nd = (n + 127) // 128
    pad = np.full(nd * 128, -1, dtype=np.int64)
    pad[:n] = sym
    blk = pad.reshape(nd, 4, 32)
    bit = (np.uint64(1) << np.arange(32, dtype=np.uint64))[None, None, :]
    def plane(mask):
        return (mask.astype(np.uint64) * bit).sum(axis=2).astype(np.uint32)
    p0 = plane((blk >= 1) & (((blk - 1) & 1) == 1))
    p1 = plane((blk >= 1) & (((blk - 1) & 2) == 2))
    p2 = plane(blk == 0)
    cnt = np.zeros((nd, 4), dtype=np.uint32)
    for c in range(1, 5):
        per = (pad.reshape(nd, 128) == c).sum(axis=1)
        cnt[:, c - 1] = np.concatenate([[0], np.cumsum(per)[:-1]]).astype(np.uint32)
    return np.concatenate([cnt, p0, p1, p2], axis=1).reshape(-1).view(np.uint8)


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid", "deep"])
def test_dense_rank_blocks_change_no_answer(name, datasets, engine):
    case, P = datasets.case(name), datasets.port(name)
    kw = dict(num_strings=case.num_strings, num_symbols=case.num_symbols)
    with engine.GpuBWT(runs=case.runs, **kw) as B, engine.GpuBWT(runs=case.runs, **kw) as D:
        B.set_dense(False)
        D.set_dense(True)
        assert D.info.dense_blocks == 1 and B.info.dense_blocks == 0
        assert np.array_equal(D.export(8), _expected_dense(case.runs))
        B.set_sai(case.sai)
        D.set_sai(case.sai)
        for tail in (True, False):
            B.set_tail_emulation(tail)
            D.set_tail_emulation(tail)
            assert D.info.dense_blocks == 1
            assert np.array_equal(B.export(7), D.export(7))          # the suffix tables agree
            idx = np.concatenate([common.positions(P.bwlen), np.array([U64MAX], dtype=np.uint64)])
            for c in b"$ACGTN\xff":
                s = np.full(idx.size, c, dtype=np.uint8)
                got = D.get_occ(s, idx)
                assert np.array_equal(got, B.get_occ(s, idx)), (chr(c), tail)
                if tail:
                    assert np.array_equal(got, P.get_occ_batch(s, idx)), chr(c)
            assert np.array_equal(D.get_char(idx[:-1]), B.get_char(idx[:-1]))
            for k in (1, 2, 12, 31, 73):
                q = case.queries(20000 if k == 31 else 3000, k, seed=40 + k)
                if k <= 31:
                    q[::97, k // 2] = ord("N")                       # flagged queries take the stray-byte path
                a, b = D.find_interval(q, k), B.find_interval(q, k)
                assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (k, tail)
                if tail:
                    lo2, up2 = P.find_interval_batch(q, k, nthreads=4)
                    assert np.array_equal(a[0], lo2) and np.array_equal(a[1], up2), k
                for strand in (0, 1):
                    assert np.array_equal(D.count(q, k, strand), B.count(q, k, strand))
            q = case.queries(3000, 31, seed=77)
            lo, up = D.find_interval(q, 31)
            hit = np.nonzero(up >= lo)[0][:400]
            if hit.size:
                rows = np.concatenate([np.arange(lo[i], min(up[i], lo[i] + 20) + 1, dtype=np.uint64) for i in hit])
                a = D.extract_prefix(rows, max_steps=128, want_read_index=True, want_bases=True)
                b = B.extract_prefix(rows, max_steps=128, want_read_index=True, want_bases=True)
                assert all(np.array_equal(x, y) for x, y in zip(a, b)), tail
                for every in (1, 8):
                    D.build_locate_table(128, every=every)
                    c = D.extract_prefix(rows, max_steps=128, want_read_index=True)
                    assert all(np.array_equal(x, y) for x, y in zip(c, b[:3])), (tail, every)
                    D.drop_locate_table()


def test_dense_blocks_on_degenerate_shapes_and_auto_policy(engine, monkeypatch):
    rng = np.random.default_rng(5)
    shapes = {
        "ones": ((rng.integers(0, 5, 20000) << 5) | 1).astype(np.uint8),            # runs of 1: auto builds dense blocks
        "single": np.array([(2 << 5) | 17], dtype=np.uint8),
        "exact128": ((rng.integers(1, 5, 64) << 5) | 2).astype(np.uint8),            # n = 128: one full dense block
        "long": ((rng.integers(1, 5, 5000) << 5) | 31).astype(np.uint8),             # long runs: auto keeps run blocks only
    }
    for name, runs in shapes.items():
        n = int((runs & 31).sum())
        ns = int((runs[(runs >> 5) == 0] & 31).sum())
        with engine.GpuBWT(runs=runs, num_strings=ns, num_symbols=n) as A:
            assert A.info.dense_blocks == (0 if name == "long" else 1), name
            A.set_dense(True)
            assert np.array_equal(A.export(8), _expected_dense(runs)), name
            P = oracle.PortIndex(runs=runs, num_strings=ns, num_symbols=n)
            idx = np.arange(n, dtype=np.uint64)
            for c in b"$ACGT":
                s = np.full(idx.size, c, dtype=np.uint8)
                assert np.array_equal(A.get_occ(s, idx), P.get_occ_batch(s, idx)), (name, chr(c))
            assert np.array_equal(A.get_char(idx), P.get_char_batch(idx)), name
    monkeypatch.setenv("RS_B200_DENSE", "0")
    runs = shapes["ones"]
    with engine.GpuBWT(runs=runs, num_strings=int(((runs >> 5) == 0).sum()), num_symbols=runs.size) as A:
        assert A.info.dense_blocks == 0


# ------------------------------------------------------------------------------------------------
# round-1 advisor findings
# ------------------------------------------------------------------------------------------------
def test_fused_count_buffers_survive_the_next_begin(datasets, engine):
    """rs_comm_finish(e)'s counts stay valid until the next-but-one rs_comm_begin (three rotating buffers)."""
    import torch
    from readserver_b200 import sharding
    case = datasets.case("clean")
    Q = 5000
    q1, q2 = case.queries(Q, 31, seed=1), case.queries(Q, 31, seed=2)
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        want1, want2 = B.count(q1, 31), B.count(q2, 31)
        comm = sharding.connect_peers(engine.PeerComm(0, 0, 1, Q), "cuda:0")
        d1, d2 = torch.from_numpy(q1).cuda(), torch.from_numpy(q2).cuda()
        comm.begin(Q)
        comm.accumulate(B, d1, 31, Q)
        p1 = comm.finish()
        comm.begin(Q)                                    # step e + 1 clears the buffer of step e + 2, not step e's
        torch.cuda.synchronize()
        assert np.array_equal(sharding.counts_view(p1, Q, "cuda:0").cpu().numpy().view(np.uint64), want1)
        comm.accumulate(B, d2, 31, Q)
        p2 = comm.finish()
        torch.cuda.synchronize()
        assert p2 != p1
        assert np.array_equal(sharding.counts_view(p2, Q, "cuda:0").cpu().numpy().view(np.uint64), want2)
        assert np.array_equal(sharding.counts_view(p1, Q, "cuda:0").cpu().numpy().view(np.uint64), want1)
        comm.close()


def test_locate_tables_when_the_tail_region_holds_no_T(engine):
    """Under the reference's tail behaviour an LF step can land on row n (rank clamped, every T already counted): the
    table builders must skip it instead of writing past their arrays.  Tail region here: the last runs are A/C/G only."""
    runs, _ = random_runs(9, 60_000)
    runs = runs.copy()
    tail = runs[-6000:]
    sym = tail >> 5
    tail[sym == 4] = (np.uint8(1) << 5) | (tail[sym == 4] & 31)       # T -> A in the last 6000 runs
    n = int((runs & 31).sum())
    ns = int((runs[(runs >> 5) == 0] & 31).sum())
    P = oracle.PortIndex(runs=runs, num_strings=ns, num_symbols=n)
    with engine.GpuBWT(runs=runs, num_strings=ns, num_symbols=n) as B:
        assert B.info.ref_clamp_c < n                                  # the tail region is not empty
        rng = np.random.default_rng(1)
        rows = np.concatenate([rng.integers(0, n, 20000).astype(np.uint64), np.arange(n - 3000, n, dtype=np.uint64)])
        base = B.extract_prefix(rows, max_steps=200)
        lex, st, _, _ = P.locate_batch(rows, prefix=True, postfix=False, max_steps=200)
        assert np.array_equal(base[0], lex) and np.array_equal(base[1], st)
        for dense in (True, False):
            B.set_dense(dense)
            for every in (1, 4, 16):
                B.build_locate_table(200, every=every)
                got = B.extract_prefix(rows, max_steps=200)
                assert np.array_equal(got[0], base[0]) and np.array_equal(got[1], base[1]), (dense, every)
                B.drop_locate_table()
        B.set_sai(np.arange(ns, dtype=np.uint64))
        with engine.ShardSet() as S:                                   # the read-index table takes the same walks
            S.add(B, 0)
            S.seal(200, 1)
