#!/usr/bin/env python
"""bench.py -- 31-mer count+locate throughput on B200 (BASELINE.json metric, configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload = BASELINE.json configs[2]: 100 diploid samples over a random 64 Mbp genome (population site list 1/300,
~1 SNP/kb per haplotype), 30x of 100-bp reads per sample, de-duplicated, partitioned into the reference's 64 shards by
reversed last 3 bases (service.cpp:820, demo/permutations-3.txt), per shard a BWT + .sai + sample table
(size_of_sample = 1); 10 M 31-mers (90 % cut from the genome).

A step = ONE pass of the hot path over the whole batch: per k-mer and per shard findInterval (query.cpp:24-41), every row
of every interval located (extractPrefix's walk -> .sai -> read index, query.cpp:43-63), the hit reads' sample codes
united into the k-mer's sample set (service.cpp:816-845), per-shard answers summed / concatenated per k-mer as the
reference's front end does (server.cpp:184-197, 404-410).  Outputs per k-mer: population count, hit list
(shard, read index), carrier bitmap.

N = 1   all 64 shards on one GPU (rs_shardset / rs_pop_query_dev).
N > 1   strong scaling: the 64 shards are placed on the N GPUs by raw read count (LPT), every rank runs the whole batch
        on its shards, and the pipeline's own kernels write counts, hit lists and carrier sets into the buffers of the
        rank owning each k-mer (first four bases mod N) over NVLink peer memory (rs_popcomm).  value = Q / time.

Prints ONE JSON line on rank 0.  `--impl reference` times the reference's own CPU code (oracle/_ref: findInterval +
extractPrefix + extractPostfix per hit row, all host threads) on a bounded sample of the same batch.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "31-mer count+locate queries/s (findInterval + locate + sample set per k-mer over 64 read-suffix shards)"
UNIT = "queries/s"
BYTES_PER_RANK = 64           # one getOcc reads one 64-byte block = two 32-byte sectors
FALLBACK_HBM_GBS = 6650.0     # /opt/skills/guides/B200_PROFILING.md
NSHARDS = 64
MAX_HITS = 1 << 20            # rows reported per k-mer and shard (the reference's front end refuses > 100,000 reads)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # workload (defaults = BASELINE.json configs[2]); --scale shrinks genome and batch together for debugging
    ap.add_argument("--genome", type=int, default=64_000_000)
    ap.add_argument("--samples", type=int, default=100)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--read-len", type=int, default=100)
    ap.add_argument("--queries", type=int, default=10_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--locate-every", type=int, default=1,
                    help="1: read-index tables (4 B per BWT row); S > 1: sampled locate tables; 0: plain walks")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side", action="store_true", help="skip the side measurements (configs[1] count line, roofline probe)")
    ap.add_argument("--c3-symbols", type=int, default=9_000_000_000, help="configs3: symbols per shard")
    ap.add_argument("--c3-shards", type=int, default=NSHARDS, help="configs3: shards in the collection (64; fewer for trials)")
    ap.add_argument("--workload", default="configs2", choices=["configs2", "configs4", "configs1", "configs3"],
                    help="configs2 (default, the metric's): 10 M 31-mers, forward strand.  configs4: variant-allele probes over "
                         "the same population -- --sites sites of its list, the reference- and the alternative-allele 31-mer "
                         "centred on each, both strands -> read indices -> carrier sets")
    ap.add_argument("--sites", type=int, default=1_000_000)
    a = ap.parse_args()
    if a.scale != 1.0:
        a.genome = max(200_000, int(a.genome * a.scale))
        a.queries = max(10_000, int(a.queries * a.scale))
        a.sites = max(5_000, int(a.sites * a.scale))
    if a.workload == "configs4":
        a.queries = 2 * a.sites
    a.strands = (0, 1) if a.workload == "configs4" else (0,)
    return a


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def workload_name(args):
    if args.workload == "configs4":
        return (f"configs[4]: {args.sites} sites of the population's list, reference- and alternative-allele {args.k}-mer centred "
                f"on each ({args.queries} k-mers), both strands -> matching read indices -> per-k-mer carrier sets; population = "
                f"configs[2]'s ({args.samples} diploid samples x {args.coverage:g}x of {args.read_len}-bp reads, random "
                f"{args.genome / 1e6:g} Mbp genome, {NSHARDS} shards by reversed last 3 bases)")
    return (f"configs[2]: {args.samples} diploid samples x {args.coverage:g}x of {args.read_len}-bp reads over a random "
            f"{args.genome / 1e6:g} Mbp genome (sites 1/300, ~1 SNP/kb per haplotype), de-duplicated, {NSHARDS} shards by "
            f"reversed last 3 bases, per-shard RLBWT + .sai + sample table; {args.queries} random {args.k}-mers (90% present), "
            f"count + locate + sample set")


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def build_population(args, device, shards_wanted=None, keep_host=False, make_index=True, local_rank=0):
    """Generates the population and builds the wanted shards (None = all).  Returns dict with per-shard entries:
    gpu (engine.GpuBWT with .sai and sample table attached) and/or host arrays for the CPU legs."""
    import torch
    from readserver_b200 import datagen as dg
    t0 = time.time()
    pop = dg.make_population(args.genome, args.samples, seed=4, device=device)
    keys, hist = dg.population_read_keys(pop, args.coverage, args.read_len, seed=41)
    hist = [int(x) for x in hist.cpu().numpy()]
    t_keys = time.time() - t0
    wanted = list(range(NSHARDS)) if shards_wanted is None else shards_wanted(hist)
    if make_index:
        from readserver_b200 import engine
    out = {"hist": hist, "shards": {}, "wanted": wanted}
    tot = dict(reads=0, symbols=0, runs=0, codes=0, dropped=0)
    for s in wanted:
        sh = dg.build_population_shard(pop, keys, s, args.read_len)
        b = sh.built
        ent = {"num_strings": b.num_strings, "num_symbols": b.num_symbols, "num_runs": int(b.runs.numel())}
        if make_index:
            B = engine.GpuBWT(runs=b.runs.contiguous(), num_strings=b.num_strings, num_symbols=b.num_symbols, device=local_rank)
            B.set_sai(b.sai.contiguous())
            B.set_samples(sh.samp_off.contiguous(), sh.samp_codes.contiguous(), 1)
            ent["gpu"] = B
        if keep_host:
            ent["runs"] = b.runs.cpu().numpy()
            ent["sai"] = b.sai.cpu().numpy().astype(np.uint64)
            ent["samp_off"] = sh.samp_off.cpu().numpy()
            ent["samp_codes"] = sh.samp_codes.cpu().numpy()
        tot["reads"] += b.num_strings
        tot["symbols"] += b.num_symbols
        tot["runs"] += int(b.runs.numel())
        tot["codes"] += int(sh.samp_codes.numel())
        tot["dropped"] += sh.dropped_reads
        out["shards"][s] = ent
        del sh, b
    del keys
    if args.workload == "configs4":
        ref, alt, _ = dg.site_probe_kmers(pop, args.sites, args.k, seed=6)
        queries = torch.stack([ref, alt], dim=1).reshape(-1, args.k).contiguous()   # ref, alt, ref, alt, ...
    else:
        queries = dg.population_queries(pop, args.queries, args.k, seed=5).contiguous()
    out.update(queries=queries, totals=tot, build_seconds=time.time() - t0, key_seconds=t_keys)
    del pop
    if device != "cpu":
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            import torch
            uuid = str(torch.cuda.get_device_properties(index).uuid)
            h = None
            for i in range(pynvml.nvmlDeviceGetCount()):
                hh = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(hh)
                u = u.decode() if isinstance(u, bytes) else u
                if uuid in u:
                    h = hh
            self.h = h if h is not None else pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.005)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU legs (the ONLY place bench.py touches oracle/)
# ------------------------------------------------------------------------------------------------
class CpuShards:
    """The shards loaded into the reference's own CPU implementation (oracle/_ref when built, else the oracle port)."""

    def __init__(self, data, read_len):
        import oracle
        import torch
        from readserver_b200 import datagen as dg
        self.ix, self.sai, self.samp = {}, {}, {}
        self.max_steps = read_len + 28
        t0 = time.time()
        with tempfile.TemporaryDirectory() as d:
            for s, ent in data["shards"].items():
                path = os.path.join(d, f"shard{s}.bwt")
                dg.write_bwt(path, torch.from_numpy(ent["runs"]), ent["num_strings"], ent["num_symbols"])
                self.ix[s] = oracle.open_best(path)
                os.unlink(path)
                self.sai[s] = ent["sai"]
                self.samp[s] = (ent["samp_off"], ent["samp_codes"])
        self.kind = next(iter(self.ix.values())).kind
        self.load_seconds = time.time() - t0

    def run(self, q, k, nthreads, keep=False):
        """count + locate of the k-mers q [n, k] in every shard, as the workers do it: findInterval, then extractPrefix +
        extractPostfix per hit row (find_reads, service.cpp:714-751).  Returns (seconds inside the reference's code,
        per-shard results when keep)."""
        secs, per = 0.0, {}
        for s, ix in self.ix.items():
            lo, up = ix.find_interval_batch(q, k, nthreads=nthreads)
            secs += ix.last_seconds
            c = np.where(up >= lo, np.minimum(up - lo + 1, MAX_HITS), 0).astype(np.int64)
            tot = int(c.sum())
            if tot:
                starts = np.repeat(lo, c)
                within = np.arange(tot, dtype=np.int64) - np.repeat(np.cumsum(c) - c, c)
                rows = (starts + within.astype(np.uint64)).astype(np.uint64)
                lex, _, ri, _ = ix.locate_batch(rows, nthreads=nthreads, prefix=True, postfix=True, max_steps=self.max_steps)
                secs += ix.last_seconds
            else:
                lex = ri = np.zeros(0, dtype=np.uint64)
            if keep:
                per[s] = (lo, up, c, lex, ri)
        return secs, per

    def expected(self, per, order, n):
        """counts, hit_offsets, hits (shard << 40 | read index, the set's order), carriers [n, 4] from per-shard results."""
        U64MAX = np.uint64(0xFFFFFFFFFFFFFFFF)
        counts = np.zeros(n, dtype=np.uint64)
        tot_h = np.zeros(n, dtype=np.int64)
        for s in order:
            lo, up, c, _, _ = per[s]
            counts += np.where(up >= lo, up - lo + 1, 0).astype(np.uint64)
            tot_h += c
        off = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(tot_h, out=off[1:])
        hits = np.zeros(int(off[n]), dtype=np.uint64)
        carr = np.zeros((n, 4), dtype=np.uint32)
        run = off[:-1].copy()
        postfix_agrees = {"hits_differing": 0, "shards_differing": []}
        for s in order:
            lo, up, c, lex, ri_post = per[s]
            tot = int(c.sum())
            if tot:
                sai = self.sai[s]
                ns = sai.size
                ri = np.where(lex < ns, sai[np.minimum(lex, ns - 1).astype(np.int64)], U64MAX)
                ndiff = int((ri != ri_post).sum())
                if ndiff:
                    postfix_agrees["hits_differing"] += ndiff
                    postfix_agrees["shards_differing"].append(int(s))
                qid = np.repeat(np.arange(n), c)
                within = np.arange(tot, dtype=np.int64) - np.repeat(np.cumsum(c) - c, c)
                hits[run[qid] + within] = np.where(ri == U64MAX, U64MAX, (np.uint64(s) << np.uint64(40)) | ri)
                # sample sets: union of (code - 34) over the codes of every hit read
                so, sc = self.samp[s]
                ok = ri != U64MAX
                r_ok, q_ok = ri[ok].astype(np.int64), qid[ok]
                ncodes = (so[r_ok + 1] - so[r_ok]).astype(np.int64)
                tc = int(ncodes.sum())
                if tc:
                    src = np.repeat(so[r_ok].astype(np.int64), ncodes) + (np.arange(tc, dtype=np.int64) - np.repeat(np.cumsum(ncodes) - ncodes, ncodes))
                    smp = sc[src].astype(np.int64) - 34
                    qq = np.repeat(q_ok, ncodes)
                    np.bitwise_or.at(carr, (qq, smp >> 5), (np.uint32(1) << (smp & 31).astype(np.uint32)))
            run += c
        return counts, off.astype(np.uint64), hits, carr, postfix_agrees

    def close(self):
        for ix in self.ix.values():
            ix.close()


def run_reference(args):
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    import torch
    device = "cuda:0" if torch.cuda.is_available() else "cpu"
    data = build_population(args, device, keep_host=True, make_index=False)
    q = data["queries"].cpu().numpy()
    cores = os.cpu_count() or 1
    cpu = CpuShards(data, args.read_len)
    probe = min(len(q), 2000)
    t, _ = cpu.run(q[:probe], args.k, cores)
    rate = probe / max(t, 1e-9)
    n = int(min(len(q), max(probe, rate * 90.0 / max(1, args.steps + args.warmup))))
    sample = np.ascontiguousarray(q[:n])
    for _ in range(args.warmup):
        cpu.run(sample, args.k, cores)
    secs = [cpu.run(sample, args.k, cores)[0] for _ in range(args.steps)]
    value = n * len(secs) / sum(secs)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(args), "queries_per_step": n, "shards": NSHARDS,
                   "bwt_symbols": data["totals"]["symbols"], "bwt_runs": data["totals"]["runs"],
                   "reads_unique": data["totals"]["reads"],
                   "note": "findInterval in all 64 shards + extractPrefix + extractPostfix per hit row (find_reads, "
                           "service.cpp:714-751); the RocksDB lookups of the sample store are not timed (no RocksDB here)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": cpu.kind,
                         "sample": f"first {n} of {len(q)} k-mers per step, all {NSHARDS} shards, {cores} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    cpu.close()
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, streaming copy)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def committed_traffic(name):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/<name>), or None."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", name)))
        return float(d["dram_bytes_per_launch"]), d.get("source")
    except Exception:
        return None, None


def pinned_array(engine, nbytes, dtype, shape):
    import ctypes as C
    p = C.c_void_p()
    rc = engine.lib().rs_host_alloc(C.byref(p), max(nbytes, 8))
    assert rc == 0
    buf = (C.c_uint8 * max(nbytes, 8)).from_address(p.value)
    return np.frombuffer(buf, dtype=dtype)[: int(np.prod(shape))].reshape(shape), p


def side_configs1(args, engine, dev, local_rank, flush):
    """configs[1] count line (round 1's headline), kept as a side entry: one 1.08 G-symbol RLBWT, 1 M 31-mers, K1 alone."""
    import torch
    from readserver_b200 import datagen as dg
    t0 = time.time()
    h1 = dg.random_genome(10_000_000, seed=2, device=dev)
    h2 = dg.mutate_snps(h1, 1e-3, seed=20)
    reads, lens = dg.sample_reads([h1, h2], 50_000_000, 100, seed=21)
    reads, lens, _, _ = dg.dedupe_rlo(reads, lens)
    built = dg.build_bwt(reads, lens)
    del reads, lens, h2
    Q, k = 1_000_000, 31
    dq = dg.make_queries(h1, Q, k, seed=3).contiguous()
    B = engine.GpuBWT(runs=built.runs.contiguous(), num_strings=built.num_strings, num_symbols=built.num_symbols, device=local_rank)
    B.enable_stats(True)
    nsym, nruns = built.num_symbols, int(built.runs.numel())
    del built
    torch.cuda.empty_cache()
    build_s = time.time() - t0
    dlo = torch.empty(Q, dtype=torch.int64, device=dev)
    dup = torch.empty(Q, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        B.find_interval_dev(dq, k, Q, dlo, dup, stream)
    torch.cuda.synchronize()
    B.read_stats(reset=True)
    ms = []
    for _ in range(20):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        B.find_interval_dev(dq, k, Q, dlo, dup, stream)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    steps, _ = B.read_stats(reset=True)
    t = sum(ms) / 1e3
    # e2e through the host-pointer call, pinned buffers
    hq, _p1 = pinned_array(engine, Q * k, np.uint8, (Q, k))
    hlo, _p2 = pinned_array(engine, Q * 8, np.uint64, (Q,))
    hup, _p3 = pinned_array(engine, Q * 8, np.uint64, (Q,))
    hq[:] = dq.cpu().numpy()
    L = engine.lib()
    for _ in range(3):
        assert L.rs_find_interval_batch(B._h, hq.ctypes.data, k, Q, hlo.ctypes.data, hup.ctypes.data) == 0
    es = 0.0
    for _ in range(10):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        assert L.rs_find_interval_batch(B._h, hq.ctypes.data, k, Q, hlo.ctypes.data, hup.ctypes.data) == 0
        es += time.perf_counter() - t0
    same = bool(np.array_equal(dlo.cpu().numpy().view(np.uint64), hlo) and np.array_equal(dup.cpu().numpy().view(np.uint64), hup))
    out = {"workload": "configs[1]: 50M x 100bp reads, random 10 Mbp diploid genome, de-duplicated, one RLBWT; 1M 31-mers",
           "metric": "31-mer count queries/s (findInterval)", "value": Q * len(ms) / t, "ms_per_step": 1e3 * t / len(ms),
           "e2e_value": Q * 10 / es, "e2e_api": "rs_find_interval_batch (host pointers, pinned, 31 B in + 16 B out per query)",
           "achieved_GBps_at_64B_per_getocc": 2.0 * steps / len(ms) * BYTES_PER_RANK / (t / len(ms)) / 1e9,
           "index_hbm_bytes": int(B.info.hbm_bytes), "bwt_symbols": nsym, "bwt_runs": nruns, "host_equals_device": same,
           "build_seconds": round(build_s, 1)}
    B.close()
    for p in (_p1, _p2, _p3):
        engine.lib().rs_host_free(p)
    del dq, dlo, dup
    torch.cuda.empty_cache()
    return out


def revcomp_rows(q: np.ndarray) -> np.ndarray:
    comp = np.zeros(256, dtype=np.uint8)
    for x, y in zip(b"ACGT", b"TGCA"):
        comp[x] = y
    return np.ascontiguousarray(comp[q[:, ::-1]])


def run_ours(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    from readserver_b200 import engine, sharding

    want_cpu = (world == 1 and not args.no_cpu_baseline)
    place = (lambda hist: sharding.assign_shards(hist, world)[rank]) if world > 1 else None
    data = build_population(args, dev, shards_wanted=place, keep_host=want_cpu, make_index=True, local_rank=local_rank)
    my_shards = data["wanted"]
    t0 = time.time()
    S = engine.ShardSet(local_rank)
    for s in my_shards:
        S.add(data["shards"][s]["gpu"], s)
    info = S.seal(args.read_len + 28, args.locate_every)
    seal_s = time.time() - t0
    base_index_bytes = int(info.index_bytes)
    S.enable_stats(True)
    Q, k, strands = args.queries, args.k, args.strands
    dq = data["queries"]
    W = int(info.carrier_words)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()  # `ncu --profile-from-start off` skips the synthetic-data build
    stream = torch.cuda.current_stream().cuda_stream
    # sizing pass per strand: hits of the batch in this rank's shards
    tmp_c = torch.empty(Q, dtype=torch.int64, device=dev)
    tmp_o = torch.empty(Q + 1, dtype=torch.int64, device=dev)
    d_total = torch.zeros(1, dtype=torch.int64, device=dev)
    local_hits = {}
    for sd in strands:
        S.query_dev(dq, k, Q, False, MAX_HITS, tmp_c, tmp_o, None, 0, None, d_total, stream, strand=sd)
        torch.cuda.synchronize()
        local_hits[sd] = int(d_total.item())
    del tmp_c, tmp_o
    out, comms, caps, total_hits = {}, {}, {}, {}
    for sd in strands:
        if world > 1:
            th = torch.tensor([local_hits[sd]], dtype=torch.int64, device=dev)
            dist.all_reduce(th)
            total_hits[sd] = int(th.item())
            caps[sd] = int(total_hits[sd] / world * 1.3) + (1 << 20)   # ownership by 4-base prefix is even for random k-mers
            comms[sd] = sharding.connect_peers(engine.PopComm(local_rank, rank, world, Q, caps[sd]), dev)
        else:
            total_hits[sd] = local_hits[sd]
            caps[sd] = max(total_hits[sd], 1)
            out[sd] = dict(counts=torch.empty(Q, dtype=torch.int64, device=dev), off=torch.empty(Q + 1, dtype=torch.int64, device=dev),
                           hits=torch.empty(caps[sd], dtype=torch.int64, device=dev),
                           carr=torch.empty((Q, W), dtype=torch.int32, device=dev))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        for sd in strands:
            if world > 1:
                comms[sd].query_dev(S, dq, k, Q, False, MAX_HITS, True, True, stream, strand=sd)
            else:
                o = out[sd]
                S.query_dev(dq, k, Q, False, MAX_HITS, o["counts"], o["off"], o["hits"], caps[sd], o["carr"], d_total, stream, strand=sd)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        assert not any(c.timed_out() for c in comms.values()), "peer barrier timed out"
    S.read_stats(reset=True)
    S.enable_timing(True)
    clocks = ClockSampler(local_rank)
    clocks.start()
    torch.cuda.synchronize()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1)  # evict queries, scratch and whatever index lines L2 still holds (untimed)
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier()  # ranks start the step together: a step's time is the step, not the skew of the flush
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clk = clocks.stop()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    stage_ms, stage_launches = S.read_timing()
    S.enable_timing(False)
    interval_steps, walk_steps, block_loads = S.read_stats(reset=True)
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    stage_by_rank = None
    if world > 1:  # every rank's stage times: the step ends at the slowest rank's last barrier
        per = [None] * world
        dist.all_gather_object(per, {s_: round(v / args.steps, 3) for s_, v in stage_ms.items()})
        stage_by_rank = per

    # ---- N > 1: the fanned-in answers against the ranks' own answers, gathered with torch.distributed ----
    fan_in_ok = None
    if world > 1:
        ok = True
        dq_np = dq.cpu().numpy()
        for sd in strands:
            comm = comms[sd]
            assert not comm.timed_out(), "peer barrier timed out"
            res = comm.results(Q, dev)
            searched = revcomp_rows(dq_np) if sd else dq_np          # ownership goes by the searched string
            owner = torch.from_numpy(sharding.owner_of(searched, world)).to(dev)
            mine = owner == rank
            lc = torch.empty(Q, dtype=torch.int64, device=dev)
            lo_ = torch.empty(Q + 1, dtype=torch.int64, device=dev)
            lh = torch.empty(max(local_hits[sd], 1), dtype=torch.int64, device=dev)
            lcar = torch.empty((Q, W), dtype=torch.int32, device=dev)
            S.query_dev(dq, k, Q, False, MAX_HITS, lc, lo_, lh, max(local_hits[sd], 1), lcar, d_total, stream, strand=sd)
            torch.cuda.synchronize()
            gc = lc.clone()
            dist.all_reduce(gc)
            ok &= bool(torch.equal(res["counts"][mine], gc[mine]) and int(res["counts"][~mine].abs().sum()) == 0)
            # hit lists: per owned k-mer the number of hits = sum of the ranks' numbers; on a sample the lists themselves
            nh = (lo_[1:] - lo_[:-1]).clone()
            dist.all_reduce(nh)
            got_n = res["hit_offsets"][1:] - res["hit_offsets"][:-1]
            ok &= bool(torch.equal(got_n[mine], nh[mine]) and int(got_n[~mine].sum()) == 0)
            ok &= int(res["total"].item()) <= caps[sd]
            nsamp = min(Q, 20000)
            h_end = int(lo_[nsamp].item())
            pieces = [None] * world
            dist.all_gather_object(pieces, (lo_[: nsamp + 1].cpu().numpy(), lh[:h_end].cpu().numpy(), lcar[:nsamp].cpu().numpy()))
            ro = res["hit_offsets"][: nsamp + 1].cpu().numpy()
            rh = res["hits"][: int(ro[-1])].cpu().numpy() if int(ro[-1]) else np.zeros(0, np.int64)
            rc = res["carriers"][:nsamp].cpu().numpy()
            own_np = owner[:nsamp].cpu().numpy()
            for i in np.nonzero(own_np == rank)[0][:4000]:
                want = np.concatenate([p[1][int(p[0][i]): int(p[0][i + 1])] for p in pieces])
                ok &= bool(np.array_equal(rh[int(ro[i]): int(ro[i + 1])], want))
                wc = np.bitwise_or.reduce(np.stack([p[2][i] for p in pieces]), axis=0)
                ok &= bool(np.array_equal(rc[i], wc))
            del lc, lo_, lh, lcar, gc, nh
        okt = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        fan_in_ok = bool(okt.item())
        assert fan_in_ok, "results fanned in through peer memory differ from the ranks' own answers"

    # ---- e2e: host buffers (pinned), every copy inside the timed region ----
    L = engine.lib()
    hq, _p0 = pinned_array(engine, Q * k, np.uint8, (Q, k))
    hq[:] = dq.cpu().numpy()
    e2e_steps = max(3, args.steps // 2)
    host, keep_alive = {}, [_p0]
    for sd in strands:
        cap = caps[sd]
        dt = np.uint64 if world == 1 else np.int64
        a1, p1 = pinned_array(engine, Q * 8, dt, (Q,))
        a2, p2 = pinned_array(engine, (Q + 1) * 8, dt, (Q + 1,))
        if world == 1:   # the hit list travels as 5 bytes per hit (rs_pop_query_hits40)
            a3, p3 = pinned_array(engine, cap * 5, np.uint8, (cap * 5,))
        else:
            a3, p3 = pinned_array(engine, cap * 8, dt, (cap,))
        a4, p4 = pinned_array(engine, Q * W * 4, np.uint32 if world == 1 else np.int32, (Q, W))
        host[sd] = dict(counts=a1, off=a2, hits=a3, carr=a4)
        keep_alive += [p1, p2, p3, p4]
    tot = C.c_uint64()
    if world == 1:
        def e2e_call():
            for sd in strands:
                h = host[sd]
                rc = L.rs_pop_query_hits40(S._h, hq.ctypes.data, k, Q, 0, sd, MAX_HITS, h["counts"].ctypes.data, h["off"].ctypes.data,
                                           h["hits"].ctypes.data, caps[sd], h["carr"].ctypes.data, C.byref(tot))
                assert rc == 0, rc
        d2h = sum(Q * 8 + (Q + 1) * 8 + total_hits[sd] * 5 + Q * W * 4 + 8 for sd in strands)
    else:
        # every rank: batch in, the pipeline with its fan-in, the answers of the k-mers it owns out
        t_hq = torch.from_numpy(hq)
        dq2 = torch.empty_like(dq)
        views = {sd: {n: torch.from_numpy(a) for n, a in host[sd].items()} for sd in strands}

        def e2e_call():
            dq2.copy_(t_hq, non_blocking=True)
            for sd in strands:
                res, v = comms[sd].results(Q, dev), views[sd]
                comms[sd].query_dev(S, dq2, k, Q, False, MAX_HITS, True, True, stream, strand=sd)
                v["counts"].copy_(res["counts"], non_blocking=True)
                v["off"].copy_(res["hit_offsets"], non_blocking=True)
                v["carr"].copy_(res["carriers"], non_blocking=True)
                torch.cuda.synchronize()
                n = int(host[sd]["off"][Q])
                v["hits"][:n].copy_(res["hits"][:n], non_blocking=True)
            torch.cuda.synchronize()
        d2h = None
    for _ in range(2):
        e2e_call()
    if world > 1:
        dist.barrier()
    e2e_s = 0.0
    for _ in range(e2e_steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        e2e_call()
        e2e_s += time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        own_hits = torch.tensor([sum(int(host[sd]["off"][Q]) for sd in strands)], dtype=torch.int64, device=dev)
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(own_hits)
        d2h = world * len(strands) * (Q * 8 + (Q + 1) * 8 + Q * 16) + int(own_hits.item()) * 8
    e2e_s = float(e2e_t.item())
    if world == 1:
        # the host path and the device path give the same answers
        for sd in strands:
            h, o = host[sd], out[sd]
            assert np.array_equal(h["counts"], o["counts"].cpu().numpy().view(np.uint64))
            assert np.array_equal(h["off"], o["off"].cpu().numpy().view(np.uint64))
            assert np.array_equal(h["carr"], o["carr"].cpu().numpy().view(np.uint32))
            n = int(h["off"][Q])
            assert n == total_hits[sd]
            for a in range(0, n, 1 << 24):   # the 5-byte list, widened piece by piece, against the device path's
                b = min(n, a + (1 << 24))
                assert np.array_equal(engine.unpack_hits40(h["hits"][5 * a: 5 * b], b - a),
                                      o["hits"][a:b].cpu().numpy().view(np.uint64))

    side = None
    if world == 1 and not args.no_side:
        side = {}
        try:
            gran = {}
            for g in (64, 128):
                gran[f"{g}B"] = engine.measure_random_access(local_rank, 16 << 30, 1 << 28, g)
            side["random_access_GBps"] = gran
        except Exception as ex:
            print(f"random-access roofline measurement failed: {ex}", file=sys.stderr)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = Q * args.steps / (total_ms / 1e3)
    peak, peak_src = hbm_peak()
    k1_s = stage_ms["search"] / 1e3
    k1_launches = max(1, stage_launches["search"])
    getocc = 2.0 * interval_steps
    achieved = block_loads * BYTES_PER_RANK / max(k1_s, 1e-12) / 1e9
    traffic, traffic_src = committed_traffic("r2_k1_traffic.json") if args.workload == "configs2" and world == 1 else (None, None)
    # this library's own kernel launches inside the timed steps, as the pipeline counted them stage by stage
    # (rs_shardset_read_timing; CUB's sort and scan kernels are not included)
    ours_launches = int(sum(stage_launches.values()))
    hits_all = sum(total_hits.values())
    line = {
        "metric": METRIC if args.workload == "configs2" else
        "variant-probe 31-mers/s (ref / alt k-mer -> read indices -> carrier set, both strands, over 64 read-suffix shards)",
        "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {
            "workload": workload_name(args), "k": k, "queries_per_step": Q, "strands": len(strands), "shards": NSHARDS,
            "shards_on_rank0": len(my_shards), "hits_per_step": hits_all, "hits_per_query": hits_all / Q,
            "locate": {1: "read-index table, 4 B per BWT row (coalesced reads of an interval's rows)", 0: "LF walk to '$' (K2)"}.get(
                args.locate_every, f"sampled locate table, every {args.locate_every} rows"),
            "bwt_symbols": data["totals"]["symbols"], "bwt_runs": data["totals"]["runs"], "reads_unique": data["totals"]["reads"],
            "sample_codes": data["totals"]["codes"], "reads_dropped_for_q2": data["totals"]["dropped"],
            "index_hbm_bytes_rank0": int(info.index_bytes), "table_hbm_bytes_rank0": int(info.table_bytes),
            "chunk_queries": int(info.chunk_queries), "carrier_words": W,
            "pipeline": ("batch sorted by packed k-mer, K1 per shard, shard-major gather (one shard per launch, rows dealt out to lanes)"
                         if world == 1 and args.locate_every == 1 and Q >= (1 << 17) and Q <= int(info.chunk_queries) else
                         "K1 per shard, query-major gather (lanes per query by the rank's shard count)" if args.locate_every == 1 else
                         "K1 per shard, per-shard walks"),
            "l2": "working set (index + tables) far larger than the 126 MB L2; 256 MiB written between timed steps (untimed)",
            "timing": "CUDA events around each step on the launching stream, summed; max over ranks"
                      + ("; ranks aligned with a barrier before each step" if world > 1 else ""),
            "collective": "none" if world == 1 else
            "fused into the pipeline's kernels over NVLink peer memory (CUDA IPC): counts atomically added into the owner's "
            "buffer, hit numbers -> owner assigns slots -> read indices written into the owner's hit list by the locate "
            "kernel, sample sets OR-ed into the owner's bitmap; flag barriers between the phases; inside the timed step",
            "fan_in_equals_ranks_own_answers": fan_in_ok,
            "data_build_seconds": round(data["build_seconds"], 1), "seal_seconds": round(seal_s, 1),
        },
        "roofline": {
            "bound": "hbm", "kernel": "k_find_interval_dense (K1), one launch per shard and chunk", "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "algorithmic_bytes": "64-byte blocks K1 had to read x 64 B: one per updateInterval step whose two positions share "
                                 "a block, two otherwise (counted on the device)",
            "blocks_per_launch": block_loads / k1_launches, "getocc_per_launch": getocc / k1_launches,
            "launches_timed": k1_launches,
            "avg_launch_ms": stage_ms["search"] / k1_launches,
            "k1_share_of_step": stage_ms["search"] / max(sum(step_ms), 1e-9),
            "stage_ms_per_step": {s: v / args.steps for s, v in stage_ms.items()},
            **({"stage_ms_per_step_by_rank": stage_by_rank} if stage_by_rank else {}),
            "note": "one launch = the whole batch against ONE shard, whose rank blocks (70 MB) and suffix table stay largely "
                    "L2-resident for the launch: the achieved figure is served by L2 + HBM together, `traffic` (ncu) says how "
                    "much of it HBM carried (DESIGN.md 6)",
        },
        "e2e": {"value": Q * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": (world if world > 1 else 1) * Q * k,
                "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "api": "rs_pop_query_hits40 (host pointers, pinned): k-mers in, counts + hit offsets + hits (5 bytes each: shard "
                       "id << 32 | read index) + carrier sets out; the "
                       "answers leave range by range of the batch while the gather works on the next range" if world == 1
                else "per rank: batch H2D, rs_pop_query_comm_dev, the owned answers D2H"},
        "gpu_launches": ours_launches,
        "clocks": clk,
    }
    if want_cpu:
        cores = os.cpu_count() or 1
        cpu = CpuShards(data, args.read_len)
        probe = min(Q, 1000)
        t, _ = cpu.run(hq[:probe], k, cores)
        n = int(min(Q, max(probe, probe / max(t, 1e-9) * args.cpu_seconds / len(strands))))
        secs, checked, all_ok = 0.0, {}, True
        for sd in strands:
            sample = np.ascontiguousarray(hq[:n])
            s1, per = cpu.run(revcomp_rows(sample) if sd else sample, k, cores, keep=True)
            secs += s1
            w_counts, w_off, w_hits, w_carr, post = cpu.expected(per, my_shards, n)
            h = host[sd]
            he = int(h["off"][n])
            c = {"population_counts": bool(np.array_equal(h["counts"][:n], w_counts)),
                 "hit_lists_read_index": bool(np.array_equal(h["off"][: n + 1], w_off) and np.array_equal(engine.unpack_hits40(h["hits"], he), w_hits)),
                 "sample_sets": bool(np.array_equal(h["carr"][:n], w_carr)), "hits_checked": he,
                 "postfix_walk_vs_prefix_walk_read_index": post}
            all_ok &= c["population_counts"] and c["hit_lists_read_index"] and c["sample_sets"]
            checked["forward" if sd == 0 else "reverse_complement"] = c
        line["cpu_baseline"] = {
            "value": n / secs, "unit": UNIT, "cores": cores, "kind": cpu.kind,
            "sample": f"first {n} of {Q} k-mers, all {NSHARDS} shards, {len(strands)} strand(s), {cores} threads striding each batch",
            "what": "findInterval per shard + extractPrefix + extractPostfix per hit row (find_reads, service.cpp:714-751); the "
                    "sample-store lookups (RocksDB) are not in the timed part",
            "load_seconds": round(cpu.load_seconds, 1),
            "gpu_results_bit_exact_on_sample": all_ok, "checked": checked,
            "note": "read index = .sai[terminal row of extractPrefix's walk]; the reference's tail behaviour (SURVEY.md Q1) can "
                    "send that walk elsewhere than extractPostfix's (select is unaffected) in shards whose tail region is "
                    "not empty: both sides reproduce it, postfix_walk_vs_prefix_walk_read_index counts such hits",
        }
        cpu.close()
        assert all_ok, "GPU results differ from the CPU reference"
    if side is not None:
        # the locate stage three ways on the first 4 M k-mers: read-index tables (the headline's), the sampled locate
        # table (every 16th row of each read's walk), plain LF walks to the '$' (K2) -- the paths for shards too large for
        # 4 bytes per row; answers compared with the tables'
        try:
            Qs = min(Q, 4_000_000)
            o = out[0]
            modes = {}
            ref_hits = None
            for every, name in ((1, "read_index_table"), (16, "sampled_every_16"), (0, "walk_to_terminator")):
                t0 = time.time()
                if every != args.locate_every or name != "read_index_table":
                    S.seal(args.read_len + 28, every)
                seal_t = time.time() - t0
                inf = S.info
                S.query_dev(dq, k, Qs, False, MAX_HITS, o["counts"], o["off"], o["hits"], caps[0], o["carr"], d_total, stream)
                torch.cuda.synchronize()
                nh_ = int(d_total.item())
                S.read_stats(reset=True)
                S.enable_timing(True)
                for _ in range(3):
                    flush.fill_(1)
                    S.query_dev(dq, k, Qs, False, MAX_HITS, o["counts"], o["off"], o["hits"], caps[0], o["carr"], d_total, stream)
                ms_, _ = S.read_timing()
                S.enable_timing(False)
                _, wsteps, _ = S.read_stats(reset=True)
                got = o["hits"][:nh_].clone()
                same = True if ref_hits is None else bool(torch.equal(got, ref_hits))
                if ref_hits is None:
                    ref_hits = got
                modes[name] = {"locate_ms": ms_["locate"] / 3, "hit_rows_per_s": nh_ / (ms_["locate"] / 3e3), "hits": nh_,
                               "lf_steps_per_hit": wsteps / 3 / max(nh_, 1), "table_bytes": int(inf.table_bytes) if every == 1 else
                               int(inf.index_bytes) - base_index_bytes, "seal_seconds": round(seal_t, 2),
                               "equals_read_index_table": same}
            side["locate_modes_first_4M_kmers"] = modes
            del ref_hits, got
        except Exception as ex:
            print(f"locate-mode side measurement failed: {ex}", file=sys.stderr)
        # the population's shards leave HBM next: building configs[1] (a 1.08 G-symbol suffix sort) needs the room
        S.close()
        for ent in data["shards"].values():
            if "gpu" in ent:
                ent.pop("gpu").close()
        out.clear()
        torch.cuda.empty_cache()
        try:
            side["configs1_count"] = side_configs1(args, engine, dev, local_rank, flush)
        except Exception as ex:
            print(f"configs[1] side measurement failed: {ex}", file=sys.stderr)
        line["side"] = side
        ra = side.get("random_access_GBps", {}).get("64B")
        if ra:
            line["roofline"]["random_access_peak_64B"] = ra
            line["roofline"]["frac_of_random_access_peak"] = achieved / ra
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = sys.stdout


# ------------------------------------------------------------------------------------------------
# configs[3]: a run-level collection far larger than anything derivable from reads here
# ------------------------------------------------------------------------------------------------
def run_configs3(args):
    """BASELINE.json configs[3] (SURVEY.md 8(d) "Config 4"): 64 shards drawn at the run level -- run symbol and length
    following the table measured on a real configs[2]-style shard -- of --c3-symbols symbols each, 64 / N shards per GPU
    (9e9 symbols = 9.6 GB of run blocks + rank blocks + tables per shard: 77 GB per GPU at N = 8); count + locate (LF
    walks) of --queries 31-mers, 90 % of them read off the shards' own LF walks, answers fanned in to the k-mers'
    owners over NVLink peer memory as in configs[2].  Parity: a slice shard of 1 % the size, same generator, loaded into
    the reference's CPU code (oracle/_ref) and into the engine on rank 0."""
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    from readserver_b200 import datagen as dg, engine, sharding
    Q, k = args.queries, args.k
    max_steps = args.read_len + 28
    nshards = args.c3_shards

    # the run table: one shard of a small population with configs[2]'s samples, coverage and read length
    t0 = time.time()
    pop = dg.make_population(2_000_000, args.samples, seed=4, device=dev)
    keys, _ = dg.population_read_keys(pop, args.coverage, args.read_len, seed=41)
    real = dg.build_population_shard(pop, keys, 0, args.read_len).built
    table = dg.run_table(real.runs).cpu()
    mean_run = float(real.num_symbols) / max(1, real.runs.numel())
    del pop, keys, real
    torch.cuda.empty_cache()

    # this rank's shards
    my_shards = [s for s in range(nshards) if s % world == rank]
    gpu, sizes = {}, {}
    for s in my_shards:
        runs, nstr, nsym = dg.runlevel_shard(table, args.c3_symbols, seed=6000 + s, device=dev)
        B = engine.GpuBWT(runs=runs, num_strings=nstr, num_symbols=nsym, device=local_rank)
        B.set_sai(torch.arange(nstr, dtype=torch.int64, device=dev))
        sizes[s] = dict(symbols=nsym, runs=int(runs.numel()), strings=nstr, hbm_bytes=int(B.info.hbm_bytes))
        gpu[s] = B
        del runs
        torch.cuda.empty_cache()
    torch.cuda.synchronize()
    load_s = time.time() - t0

    # queries: 90 % read off LF walks from random rows of the shards (every rank contributes its share), 10 % random
    stream = torch.cuda.current_stream().cuda_stream
    share = (int(0.9 * Q) + world - 1) // world
    got = []
    g = torch.Generator(device=dev)
    g.manual_seed(500 + rank)
    need, walk_len = share, 200
    while need > 0:
        for s in my_shards:
            m = int(need / len(my_shards) / 0.55) + 1024
            rows = torch.randint(0, sizes[s]["symbols"], (m,), generator=g, device=dev, dtype=torch.int64)
            bases = torch.zeros((m, walk_len), dtype=torch.uint8, device=dev)
            steps = torch.zeros(m, dtype=torch.int32, device=dev)
            gpu[s].extract_prefix_dev(rows, m, walk_len, None, steps, None, bases, stream)
            torch.cuda.synchronize()
            ok = (steps >= k) & (steps <= walk_len)
            idx = torch.nonzero(ok).squeeze(1)
            st = steps[idx].to(torch.int64)
            cols = (st - k).unsqueeze(1) + torch.arange(k, device=dev).unsqueeze(0)   # the k characters next to the row
            got.append(torch.gather(bases[idx], 1, cols))
            del rows, bases, steps
        have = sum(int(x.shape[0]) for x in got)
        need = share - have
    present = torch.cat(got)[:share].contiguous()
    del got
    if world > 1:
        parts = [torch.empty_like(present) for _ in range(world)]
        dist.all_gather(parts, present)
        present = torch.cat(parts)
    present = present[: int(0.9 * Q)]
    gq = torch.Generator(device=dev)
    gq.manual_seed(77)
    rnd = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=dev)[
        torch.randint(0, 4, (Q - present.shape[0], k), generator=gq, device=dev)]
    dq = torch.cat([present, rnd])
    dq = dq[torch.randperm(Q, generator=gq, device=dev)].contiguous()   # the same batch on every rank
    del present, rnd
    torch.cuda.empty_cache()
    # shard set (plain LF walks: the locate tables of shards this size take minutes to build), sizing pass, communicator
    S = engine.ShardSet(local_rank)
    for s_ in my_shards:
        S.add(gpu[s_], s_)
    info = S.seal(max_steps, 0)
    S.enable_stats(True)
    l_counts = torch.empty(Q, dtype=torch.int64, device=dev)
    l_off = torch.empty(Q + 1, dtype=torch.int64, device=dev)
    d_total = torch.zeros(1, dtype=torch.int64, device=dev)
    S.query_dev(dq, k, Q, False, MAX_HITS, l_counts, l_off, None, 0, None, d_total, stream)
    torch.cuda.synchronize()
    local_hits = int(d_total.item())
    th = torch.tensor([local_hits], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(th)
    total_hits = int(th.item())
    comm = None
    if world > 1:
        cap = int(total_hits / world * 1.3) + (1 << 20)
        comm = sharding.connect_peers(engine.PopComm(local_rank, rank, world, Q, cap), dev)
    else:
        cap = max(total_hits, 1)
        o_counts = torch.empty(Q, dtype=torch.int64, device=dev)
        o_off = torch.empty(Q + 1, dtype=torch.int64, device=dev)
        o_hits = torch.empty(cap, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step():
        if world > 1:
            comm.query_dev(S, dq, k, Q, False, MAX_HITS, True, False, stream)
        else:
            S.query_dev(dq, k, Q, False, MAX_HITS, o_counts, o_off, o_hits, cap, None, d_total, stream)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        assert not comm.timed_out(), "peer barrier timed out"
    S.read_stats(reset=True)
    S.enable_timing(True)
    clocks = ClockSampler(local_rank)
    clocks.start()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clk = clocks.stop()
    stage_ms, stage_launches = S.read_timing()
    S.enable_timing(False)
    interval_steps, walk_steps, block_loads = S.read_stats(reset=True)
    tms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    total_ms = float(tms.item())

    # checks at full size: what the owners hold = the sum of the ranks' own answers; the k-mers read off the walks are found
    if world > 1:
        res = comm.results(Q, dev)
        own = torch.from_numpy(sharding.owner_of(dq.cpu().numpy(), world)).to(dev) == rank
        sums = l_counts.clone()
        dist.all_reduce(sums)
        hsum = (l_off[1:] - l_off[:-1]).clone()
        dist.all_reduce(hsum)
        ok = bool(torch.equal(res["counts"][:Q][own], sums[own]))
        ro = res["hit_offsets"][: Q + 1]
        ok = ok and bool(torch.equal((ro[1:] - ro[:-1])[own], hsum[own]))
        okt = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        fan_in_ok = bool(okt.item())
        found = int((sums > 0).sum().item())
        index_bytes = torch.tensor([int(info.index_bytes)], dtype=torch.int64, device=dev)
        allb = [torch.zeros_like(index_bytes) for _ in range(world)]
        dist.all_gather(allb, index_bytes)
        index_bytes_by_rank = [int(x.item()) for x in allb]
        per = [None] * world
        dist.all_gather_object(per, {s_: round(v / args.steps, 3) for s_, v in stage_ms.items()})
    else:
        fan_in_ok = None
        found = int((o_counts > 0).sum().item())
        assert torch.equal(o_counts, l_counts)
        index_bytes_by_rank = [int(info.index_bytes)]
        per = None
    assert fan_in_ok is not False, "results fanned in through peer memory differ from the ranks' own answers"
    # parity slice (rank 0): a shard of 1 % the size from the same generator, in the reference's CPU code and in the engine
    slice_line = None
    if rank == 0 and not args.no_cpu_baseline:
        import oracle
        nsl = max(2_000_000, args.c3_symbols // 100)
        runs, nstr, nsym = dg.runlevel_shard(table, nsl, seed=6999, device=dev)
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "slice.bwt")
            dg.write_bwt(path, runs.cpu(), nstr, nsym)
            ref = oracle.open_best(path)
        Bs = engine.GpuBWT(runs=runs, num_strings=nstr, num_symbols=nsym, device=local_rank)
        Bs.set_sai(torch.arange(nstr, dtype=torch.int64, device=dev))
        m = 200_000
        rows = torch.randint(0, nsym, (m,), generator=g, device=dev, dtype=torch.int64)
        bases = torch.zeros((m, 200), dtype=torch.uint8, device=dev)
        steps = torch.zeros(m, dtype=torch.int32, device=dev)
        Bs.extract_prefix_dev(rows, m, 200, None, steps, None, bases, stream)
        torch.cuda.synchronize()
        idx = torch.nonzero((steps >= k) & (steps <= 200)).squeeze(1)[:50_000]
        cols = (steps[idx].to(torch.int64) - k).unsqueeze(1) + torch.arange(k, device=dev).unsqueeze(0)
        qs = torch.cat([torch.gather(bases[idx], 1, cols), dq[:50_000]]).cpu().numpy()
        with engine.ShardSet(local_rank) as Ss:
            Ss.add(Bs, 0)
            Ss.seal(max_steps, 0)
            got = Ss.query(qs, k, max_hits=MAX_HITS)
        nthreads = os.cpu_count() or 1
        lo, up = ref.find_interval_batch(qs, k, nthreads=nthreads)
        secs = ref.last_seconds
        c = np.where(up >= lo, np.minimum(up - lo + 1, MAX_HITS), 0).astype(np.int64)
        rws = (np.repeat(lo, c) + (np.arange(int(c.sum()), dtype=np.int64) - np.repeat(np.cumsum(c) - c, c)).astype(np.uint64))
        lex, _, _, _ = ref.locate_batch(rws.astype(np.uint64), nthreads=nthreads, prefix=True, postfix=False, max_steps=max_steps)
        secs += ref.last_seconds
        want_hits = np.where(lex < nstr, lex, np.iinfo(np.uint64).max)   # shard id 0, .sai = identity
        slice_line = {
            "symbols": nsym, "runs": int(runs.numel()), "queries": int(qs.shape[0]), "hits": int(c.sum()),
            "counts_equal": bool(np.array_equal(got["counts"], c.astype(np.uint64))),
            "read_indices_equal": bool(np.array_equal(got["hits"], want_hits)),
            "reference": {"kind": ref.kind, "queries_per_s": qs.shape[0] / max(secs, 1e-9), "threads": nthreads}}
        assert slice_line["counts_equal"] and slice_line["read_indices_equal"], slice_line
        Bs.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = hbm_peak()
    k1_s = stage_ms["search"] / 1e3
    achieved = block_loads * BYTES_PER_RANK / max(k1_s, 1e-12) / 1e9
    tot_sym = sum(v["symbols"] for v in sizes.values())
    line = {
        "metric": "31-mer count+locate queries/s (findInterval + LF-walk locate over 64 run-level shards)",
        "value": Q * args.steps / (total_ms / 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {
            "workload": (f"configs[3]: {nshards} shards drawn at the run level (symbol and length of each run from the table of a "
                         f"real shard: {args.samples} diploid samples x {args.coverage:g}x of {args.read_len}-bp reads), "
                         f"{args.c3_symbols / 1e9:g} G symbols per shard, {nshards // world} shards per GPU; {Q} {k}-mers (90% read "
                         f"off the shards' LF walks), count + locate by LF walk (<= {max_steps} steps), answers fanned in to the "
                         f"k-mers' owners"),
            "k": k, "queries_per_step": Q, "shards": nshards, "shards_on_rank0": len(my_shards),
            "symbols_on_rank0": tot_sym, "runs_on_rank0": sum(v["runs"] for v in sizes.values()),
            "mean_run_length_of_the_real_shard": mean_run,
            "index_hbm_bytes_by_rank": index_bytes_by_rank, "index_hbm_bytes_total": sum(index_bytes_by_rank),
            "generate_and_load_seconds_rank0": round(load_s, 1), "hits_per_step": total_hits, "kmers_found": found,
            "chunk_queries": int(info.chunk_queries),
            "l2": "index per GPU far larger than the 126 MB L2 (a shard's rank blocks alone are 4.5 GB); 256 MiB written "
                  "between timed steps (untimed)",
            "timing": "CUDA events around each step on the launching stream, summed; max over ranks; ranks aligned with a "
                      "barrier before each step",
            "collective": "fused into the pipeline's kernels over NVLink peer memory (counts added into the owner's buffer, "
                          "hit numbers -> owner assigns slots -> read indices written into the owner's list)" if world > 1 else "none",
            "fan_in_equals_ranks_own_answers": fan_in_ok, "parity_slice": slice_line},
        "roofline": {
            "bound": "hbm", "kernel": "k_find_interval_dense (K1), one launch per shard", "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes": "64-byte rank blocks K1 had to read x 64 B (counted on the device)",
            "blocks_per_launch": block_loads / max(1, stage_launches["search"]),
            "stage_ms_per_step": {s_: v / args.steps for s_, v in stage_ms.items()},
            **({"stage_ms_per_step_by_rank": per} if per else {}),
            "lf_steps_per_step": walk_steps / args.steps},
        "e2e": None, "gpu_launches": int(sum(stage_launches.values()) / max(1, args.steps)), "clocks": clk,
        "cpu_baseline": ({"value": slice_line["reference"]["queries_per_s"], "unit": UNIT, "cores": slice_line["reference"]["threads"],
                          "kind": slice_line["reference"]["kind"],
                          "sample": f"{slice_line['queries']} k-mers, count + locate, against the parity slice shard "
                                    f"({slice_line['symbols']} symbols): ONE shard of 1 % the size, not the collection"}
                         if slice_line else None),
    }
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    # libraries (NCCL's version banner, torchrun notices) must not share stdout with the ONE JSON line
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.workload == "configs1":   # experiments only: round 1's count line alone (not the driver's contract)
        import torch
        from readserver_b200 import engine
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
        print(json.dumps(side_configs1(args, engine, "cuda:0", 0, flush)), file=_REAL_STDOUT, flush=True)
    elif args.impl == "reference":
        run_reference(args)
    elif args.workload == "configs3":
        run_configs3(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
