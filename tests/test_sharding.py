"""The N > 1 path on the CPU: world_size-2 `gloo` process group.  Reads are split into shards the way
the reference does (by their reversed last 3 bases), each rank holds half of the shards, per-shard
counts / hits come from the oracle (no GPU here) and the collectives of readserver_b200.sharding must
reproduce what the reference's front end computes: the sum over all shards, and the concatenated hits."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from readserver_b200 import sharding


def test_assign_shards_is_balanced_and_complete():
    sizes = [5, 9, 1, 7, 3, 3, 8, 2, 6, 4, 10, 1]
    for world in (1, 2, 4, 8):
        a = sharding.assign_shards(sizes, world)
        assert sorted(i for r in a for i in r) == list(range(len(sizes)))
        loads = [sum(sizes[i] for i in r) for r in a]
        assert max(loads) - min(loads) <= max(sizes)
    assert sharding.assign_shards([4] * 64, 8) == [sorted(range(r, 64, 8)) for r in range(8)] or \
        all(len(r) == 8 for r in sharding.assign_shards([4] * 64, 8))


def test_shard_keys_and_owner():
    keys = sharding.permutation_keys(3)
    assert len(keys) == 64 and keys[0] == "AAA" and keys[-1] == "TTT" and keys[1] == "AAC"
    ref = [l.strip() for l in open("/root/reference/demo/permutations-3.txt")] if os.path.exists(
        "/root/reference/demo/permutations-3.txt") else None
    if ref:
        assert sorted(keys) == sorted(k for k in ref if k)
    assert sharding.shard_of_read("ACGTTGCA") == "ACG"
    km = np.frombuffer(b"AAAAC" b"AAACT" b"TTTTG" b"NAAAC", dtype=np.uint8).reshape(4, 5)
    assert list(sharding.owner_of(km, 8, prefix_len=4)) == [0, 1, 255 % 8, 0]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from readserver_b200 import datagen as dg
        from tests import common
        # one read set, split into shards by reversed last 3 bases (service.cpp:820)
        case = common.make_case("shards", 24000, 6800, 3, 73, keep_reads=True)
        keys = sharding.permutation_keys(3)
        by_key = {k: [] for k in keys}
        for r in case.reads:
            by_key[sharding.shard_of_read(r)].append(r)
        sizes = [sum(len(r) for r in by_key[k]) for k in keys]
        mine = sharding.assign_shards(sizes, world)[rank]
        shards, exact = [], []
        for s in mine:
            reads = by_key[keys[s]]
            if not reads:
                continue
            rd, ln = dg.strings_to_reads(reads)
            rd, ln, _, _ = dg.dedupe_rlo(rd, ln)
            b = dg.build_bwt(rd, ln)
            shards.append(oracle.PortIndex(runs=b.runs.numpy(), num_strings=b.num_strings, num_symbols=b.num_symbols))
            E = common.Emul(common.Case(keys[s], b.runs.numpy().copy(), b.num_strings, b.num_symbols, None, None))
            E.set_tail(False)   # exact rank: lets the shard sum be checked against brute-force truth
            exact.append(E)
        q = case.queries(400, 31, seed=5)

        def counter(P):
            def f(kmers, strand):
                lo, up = P.find_interval_batch(kmers, 31)
                return np.where(up >= lo, up - lo + 1, 0).astype(np.int64)
            return f

        sc = sharding.ShardedCounter([counter(P) for P in shards])
        total = sc.count(q)
        # the whole read set in ONE index must give the same population counts where the reference is exact
        truth = np.array([common.brute_count(case.reads, bytes(k).decode()) for k in q])
        local = np.zeros(len(q), dtype=np.int64)
        for P in shards:
            local += counter(P)(q, 0)
        gathered = [torch.zeros(len(q), dtype=torch.int64) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(local))
        assert np.array_equal(total.numpy(), sum(g.numpy() for g in gathered))
        # per-shard BWTs are tiny (depth 1-2) so the reference's tail distorts their counts; with exact rank
        # the sum over all shards of both ranks is the number of occurrences in the whole read set
        ex = np.zeros(len(q), dtype=np.int64)
        for E in exact:
            lo, up = E.find_interval(q, 31)
            ex += np.where(up >= lo, up - lo + 1, 0).astype(np.int64)
        ex = sharding.reduce_counts(torch.from_numpy(ex)).numpy()
        assert np.array_equal(ex, truth), (rank, int((ex != truth).sum()), ex[:8].tolist(), truth[:8].tolist(), len(exact), len(shards))
        assert truth.sum() > 1000  # (the reference-semantics totals may be all zero on shards this small)
        # reduce-scatter with prefix-routed ownership returns exactly the owner's slice of the totals
        owners = torch.from_numpy(sharding.owner_of(q, world).astype(np.int64))
        part = sharding.reduce_scatter_counts(torch.from_numpy(local.copy()), owners)
        assert np.array_equal(part.numpy(), total.numpy()[owners.numpy() == rank])
        # C2: hit lists
        rows_off = np.zeros(len(q) + 1, dtype=np.int64)
        rows = []
        for i, k in enumerate(q):
            mine_hits = []
            for si, P in enumerate(shards):
                lo, up = P.find_interval(bytes(k).decode())
                if up >= lo:
                    mine_hits += [rank * 10**9 + si * 10**6 + int(x) for x in range(lo, min(up, lo + 4) + 1)]
            rows += mine_hits
            rows_off[i + 1] = len(rows)
        off, hits, src = sharding.gather_hits(torch.from_numpy(rows_off), torch.tensor(rows, dtype=torch.int64))
        all_rows = [None] * world
        dist.all_gather_object(all_rows, (rows_off.tolist(), rows))
        for i in range(len(q)):
            want = []
            for r in range(world):
                o, h = all_rows[r]
                want += h[o[i]:o[i + 1]]
            assert hits[int(off[i]):int(off[i + 1])].tolist() == want
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_world_size_2_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
