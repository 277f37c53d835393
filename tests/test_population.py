"""The population generator (BASELINE.json configs[2] shape): reads of many diploid samples,
partitioned into the reference's 64 shards by reversed last 3 bases (service.cpp:820, load_data_into_rocksdb.cpp:45),
per shard the unique reads in RLO order with their sample codes in merge order (rlosort_seq_and_convert_sample_names.cpp:
47-53, merge_rlosorted_reads.cpp:45-47).  Checked here against a naive regeneration of every read string."""
from collections import defaultdict

import numpy as np
import torch

from readserver_b200 import datagen as dg
from readserver_b200 import sharding


def naive_reads(pop, coverage, read_len, seed):
    G = pop.genome.numel()
    haps = []
    for h in range(pop.n_haplotypes):
        x = pop.genome.clone()
        x[pop.site_pos[pop.hap_alt[h]]] = pop.site_alt[pop.hap_alt[h]]
        haps.append(dg.ASCII[x.numpy()].tobytes().decode())
    per = int(coverage * G / read_len)
    out = []
    for s in range(pop.n_samples):
        g = dg._gen(seed * 1_000_003 + s, "cpu")
        start = torch.randint(0, G - read_len + 1, (per,), generator=g)
        hap = torch.randint(0, 2, (per,), generator=g) + 2 * s
        for st, h in zip(start.tolist(), hap.tolist()):
            out.append((haps[h][st:st + read_len], s))
    return out


def test_population_shards_equal_naive_regeneration():
    pop = dg.make_population(60_000, 5, seed=4, site_rate=1 / 50, snp_rate=1 / 200)   # dense sites: many patterns
    L = 60
    keys, hist = dg.population_read_keys(pop, 25, L, seed=41)
    allreads = naive_reads(pop, 25, L, 41)
    assert int(hist.sum()) == len(allreads) == keys.numel()
    names = sharding.permutation_keys(3)
    seen = 0
    for shard in (0, 7, 31, 32, 63):
        sh = dg.build_population_shard(pop, keys, shard, L)
        strs = dg.reads_to_strings(sh.reads, torch.full((sh.reads.shape[0],), L, dtype=torch.int32))
        d = defaultdict(list)
        for r, s in allreads:
            if sharding.shard_of_read(r) == names[shard]:
                d[r].append(s)
        want = sorted(d.keys(), key=lambda r: r[::-1])
        if sh.dropped_reads:
            want = want[: len(want) - sh.dropped_reads]
        assert want == strs
        off, codes = sh.samp_off.numpy(), sh.samp_codes.numpy()
        for i, r in enumerate(strs):
            assert [int(c) - 34 for c in codes[off[i]:off[i + 1]]] == sorted(d[r])
        assert sh.raw_reads == int(hist[shard])
        assert dg.q2_safe(sh.built.runs, sh.built.num_symbols)
        seen += len(strs)
    assert seen > 0


def test_only_the_requested_shards_are_kept_and_the_histogram_is_global():
    pop = dg.make_population(30_000, 3, seed=9)
    all_keys, hist = dg.population_read_keys(pop, 10, 50, seed=5)
    mine, hist2 = dg.population_read_keys(pop, 10, 50, seed=5, shards=[3, 40])
    assert torch.equal(hist, hist2)
    sel = ((all_keys >> 57) == 3) | ((all_keys >> 57) == 40)
    assert torch.equal(all_keys[sel], mine)
    a = dg.build_population_shard(pop, all_keys, 40, 50)
    b = dg.build_population_shard(pop, mine, 40, 50)
    assert torch.equal(a.built.runs, b.built.runs) and torch.equal(a.samp_codes, b.samp_codes)


def test_site_probes_differ_at_the_centre_only():
    pop = dg.make_population(20_000, 2, seed=3)
    ref, alt, pick = dg.site_probe_kmers(pop, 50, 31, seed=8)
    diff = (ref != alt).numpy()
    assert diff[:, 15].all() and diff.sum() == 50
    g = dg.ASCII[pop.genome.numpy()]
    for i in range(50):
        p = int(pop.site_pos[pick[i]])
        assert np.array_equal(ref[i].numpy(), g[p - 15:p + 16])


def test_runlevel_shard_follows_the_run_table():
    """configs[3]'s generator: run bytes drawn from a (symbol, length) table; the table of the output matches the input's
    proportions, the counts it returns are those of the bytes, and the bytes load into the oracle as a BWT."""
    import torch
    from readserver_b200 import datagen as dg
    import oracle
    t = torch.zeros(5, 32, dtype=torch.float64)
    t[0, 1], t[0, 2] = 40, 2
    for c in range(1, 5):
        for ln, w in ((1, 300), (2, 180), (3, 120), (7, 40), (31, 5)):
            t[c, ln] = w * (1 + 0.1 * c)
    runs, nstr, nsym = dg.runlevel_shard(t, 3_000_000, seed=11, device="cpu", chunk=250_000)
    r = runs.numpy()
    assert int((r & 31).min()) >= 1 and int((r >> 5).max()) <= 4
    assert nsym == int((r & 31).astype(np.int64).sum()) and abs(nsym - 3_000_000) < 30_000
    assert nstr == int((r[(r >> 5) == 0] & 31).astype(np.int64).sum()) > 0
    got = dg.run_table(runs)
    assert float(((got / got.sum() - t / t.sum()).abs()).max()) < 2e-3
    again, nstr2, nsym2 = dg.runlevel_shard(t, 3_000_000, seed=11, device="cpu", chunk=250_000)
    assert torch.equal(again, runs) and (nstr2, nsym2) == (nstr, nsym)          # deterministic per seed
    P = oracle.PortIndex(runs=r, num_strings=nstr, num_symbols=nsym)              # a BWT the oracle accepts
    assert P.bwlen == nsym
    lo, up = P.find_interval("ACG")
    assert up >= lo


def test_unpack_hits40_widens_the_five_byte_list():
    """engine.unpack_hits40: the 5-byte hits of rs_pop_query_hits40 (shard id << 32 | read index, little-endian; 2^40 - 1 =
    no read) in rs_pop_query's 64-bit form (shard id << 40 | read index; 2^64 - 1)."""
    from readserver_b200 import engine
    rng = np.random.default_rng(5)
    shard = rng.integers(0, 255, 1000).astype(np.uint64)
    ri = rng.integers(0, 1 << 32, 1000).astype(np.uint64)
    v = (shard << np.uint64(32)) | ri
    v[::37] = (1 << 40) - 1
    raw = np.zeros(5 * 1000 + 3, dtype=np.uint8)
    for x in range(5):
        raw[x:5000:5] = ((v >> np.uint64(8 * x)) & np.uint64(255)).astype(np.uint8)
    want = (shard << np.uint64(40)) | ri
    want[::37] = np.iinfo(np.uint64).max
    assert np.array_equal(engine.unpack_hits40(raw, 1000), want)
    assert engine.unpack_hits40(raw, 0).size == 0
