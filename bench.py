#!/usr/bin/env python
"""bench.py -- 31-mer FM-index query throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path (findInterval, src/bwt/query.cpp:24-41) over one batch of Q
synthetic 31-mers against one BWT shard per GPU.

N = 1     workload = BASELINE.json configs[1]: 50M x 100bp reads from a random 10 Mbp diploid genome,
          de-duplicated, one RLBWT; 1M 31-mers (90% cut from the genome).
N > 1     weak scaling: every rank holds its OWN shard of that size (a different read sample of the same
          genome -- the reference shards by read, and every query visits every shard: SURVEY.md 8(e)),
          all ranks search the same Q k-mers and the search kernel itself adds each per-shard count into the
          buffer of the rank owning the k-mer (by prefix) over NVLink peer memory, followed by one flag
          barrier -- inside the timed region (the reference's ZMQ fan-in + sum, server.cpp:184-197, 404-410).
          The unit is therefore one 31-mer searched in one shard; value = N*Q*K / max-over-ranks time.

Prints ONE JSON line on rank 0.  `--impl reference` times the reference's own CPU code (oracle/_ref, or
the oracle port when that library is absent) on the host cores for the same workload.
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

METRIC = "31-mer count queries/s (findInterval, one shard per GPU)"
UNIT = "queries/s"
BYTES_PER_RANK = 64           # one getOcc reads one 64-byte block = two 32-byte sectors;
                              # SURVEY.md 8(d) assumed a whole 128-byte line per getOcc (reported beside it)
FALLBACK_HBM_GBS = 6650.0     # /opt/skills/guides/B200_PROFILING.md


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # workload overrides (debugging only; the defaults are BASELINE.json configs[1])
    ap.add_argument("--genome", type=int, default=10_000_000)
    ap.add_argument("--reads", type=int, default=50_000_000)
    ap.add_argument("--read-len", type=int, default=100)
    ap.add_argument("--queries", type=int, default=1_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-random-roofline", action="store_true")
    ap.add_argument("--sharded", type=int, default=64, choices=[0, 4, 16, 64],
                    help="shards of the side measurement that splits the read set as the reference does (0 = skip)")
    ap.add_argument("--locate-queries", type=int, default=100_000, help="k-mers of the count+locate side measurement")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def build_shard(args, device, rank):
    """configs[1] for shard `rank`: (runs u8 [R] numpy, num_strings, num_symbols, sai numpy, queries [Q,k] tensor)."""
    import torch
    from readserver_b200 import datagen as dg
    t0 = time.time()
    h1 = dg.random_genome(args.genome, seed=2, device=device)
    h2 = dg.mutate_snps(h1, 1e-3, seed=20)
    for attempt in range(4):  # re-seed until the reference would not read m_vSum out of bounds
        reads, lens = dg.sample_reads([h1, h2], args.reads, args.read_len, seed=21 + 100 * rank + 7919 * attempt)
        reads, lens, _, _ = dg.dedupe_rlo(reads, lens)
        built = dg.build_bwt(reads, lens)
        if dg.q2_safe(built.runs, built.num_symbols):
            break
    queries = dg.make_queries(h1, args.queries, args.k, seed=3).contiguous()
    runs = built.runs.cpu().numpy()
    sai = built.sai.cpu().numpy().astype(np.uint64)
    ns, nsym = built.num_strings, built.num_symbols
    # the reads in builder input order, for the packed read table of the locate side measurement (rank 0 only)
    reads_text = reads_off = reads_dev = None
    if rank == 0 and args.impl == "ours" and int(lens.min()) == int(lens.max()):
        if args.locate_queries > 0:
            reads_text = dg.ASCII[reads.cpu().numpy()].reshape(-1)
            reads_off = np.arange(ns + 1, dtype=np.uint64) * np.uint64(reads.shape[1])
        if args.sharded > 0 and device != "cpu":
            reads_dev = reads
    del built, h2, reads, lens
    if device != "cpu":
        torch.cuda.empty_cache()
    return dict(runs=runs, num_strings=ns, num_symbols=nsym, sai=sai, queries=queries, genome=h1,
                reads_text=reads_text, reads_off=reads_off, reads_dev=reads_dev, build_seconds=time.time() - t0)


def workload_name(args):
    return (f"configs[1]: {args.reads // 1_000_000}M x {args.read_len}bp reads, random {args.genome / 1e6:g} Mbp diploid "
            f"genome (SNPs 1/kb), de-duplicated, single RLBWT per GPU; {args.queries} random {args.k}-mers (90% present)")


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
            # CUDA_VISIBLE_DEVICES may remap; use the UUID of the torch device
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
            self._stop.wait(0.002)

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
def cpu_find_interval(shard, queries_np, k, nthreads, target_seconds, steps=1, warmup=0, extras=False):
    """Times findInterval of the reference's CPU implementation on a bounded sample of the queries."""
    import oracle
    import torch
    from readserver_b200 import datagen as dg
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "shard.bwt")
        dg.write_bwt(path, torch.from_numpy(shard["runs"]), shard["num_strings"], shard["num_symbols"])
        t0 = time.time()
        ix = oracle.open_best(path)
        load_s = time.time() - t0
    probe = min(len(queries_np), 20_000)
    ix.find_interval_batch(queries_np[:probe], k, nthreads=nthreads)
    rate = probe / max(ix.last_seconds, 1e-9)
    n = int(min(len(queries_np), max(probe, rate * target_seconds / max(1, steps + warmup))))
    sample = np.ascontiguousarray(queries_np[:n])
    for _ in range(warmup):
        ix.find_interval_batch(sample, k, nthreads=nthreads)
    secs = []
    for _ in range(steps):
        lo, up = ix.find_interval_batch(sample, k, nthreads=nthreads)
        secs.append(ix.last_seconds)
    out = dict(kind=ix.kind, n=n, seconds=secs, value=n * len(secs) / sum(secs), load_seconds=load_s,
               lower=lo, upper=up)
    if extras and ix.kind == "reference":
        # (a) the reference as its Makefiles ship it (no -O flag), a few seconds of it; (b) the CPU locate leg:
        # extractPrefix per hit row, the walk K2 replaces
        try:
            with tempfile.TemporaryDirectory() as d:
                path = os.path.join(d, "shard.bwt")
                dg.write_bwt(path, torch.from_numpy(shard["runs"]), shard["num_strings"], shard["num_symbols"])
                slow = oracle.RefIndex(path, so=oracle.REF_O0_SO)
            m = min(n, 200_000)
            slow.find_interval_batch(sample[:m], k, nthreads=nthreads)
            out["as_shipped_O0_value"] = m / max(slow.last_seconds, 1e-9)
            slow.close()
        except Exception as ex:  # the -O0 build is optional
            out["as_shipped_O0_value"] = None
            print(f"as-shipped reference leg skipped: {ex}", file=sys.stderr)
        hit = np.nonzero(up >= lo)[0][:20_000]
        rows = np.concatenate([np.arange(lo[i], min(up[i], lo[i] + 15) + 1, dtype=np.uint64) for i in hit]) if hit.size else np.zeros(0, np.uint64)
        if rows.size:
            rows = rows[:300_000]
            ix.locate_batch(rows, nthreads=nthreads, prefix=True, postfix=False, max_steps=shard_read_len(shard) + 28)
            out["locate_rows_per_s"] = rows.size / max(ix.last_seconds, 1e-9)
            out["locate_rows"] = int(rows.size)
    return out


def shard_read_len(shard):
    return int(shard["num_symbols"] // max(1, shard["num_strings"]))


def run_reference(args):
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    import torch
    device = "cuda:0" if torch.cuda.is_available() else "cpu"
    shard = build_shard(args, device, 0)
    q = shard["queries"].cpu().numpy()
    cores = os.cpu_count() or 1
    r = cpu_find_interval(shard, q, args.k, cores, target_seconds=60.0, steps=args.steps, warmup=args.warmup)
    ms = 1e3 * sum(r["seconds"]) / len(r["seconds"])
    sample = f"first {r['n']} of {len(q)} queries per step, {cores} threads striding the array"
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(args), "bwt_symbols": shard["num_symbols"], "bwt_runs": int(shard["runs"].size),
                   "queries_per_step": r["n"]},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": cores, "kind": r["kind"], "sample": sample},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
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


def k1_traffic():
    """dram__bytes_read + dram__bytes_write of one K1 launch on this workload, from the committed ncu --set full
    capture (profiles/k1_traffic.json, written by profiles/ncu_traffic.py); None when there is none."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "k1_traffic.json")))
        return float(d["dram_bytes_per_launch"]), d.get("source")
    except Exception:
        return None, None


def sharded_measure(args, shard, B_whole, dq, engine, local_rank, stream, flush):
    """Side measurement (configs[2] shape): the read set split into 4^p shards by its reads' reversed last p bases
    (p = 3: the 64 keys of demo/permutations-3.txt, service.cpp:820), one BWT per shard, all on this GPU; every
    shard searches the whole batch and the search kernel sums the per-shard counts (the front end's fan-in,
    server.cpp:184-197, 404-410).  Checked: with exact rank the sum equals the count in the unsplit index."""
    import torch
    from readserver_b200 import datagen as dg
    from readserver_b200 import sharding
    reads = shard["reads_dev"]
    n, L = reads.shape
    p = {4: 1, 16: 2, 64: 3}[args.sharded]
    dev = reads.device
    key = torch.zeros(n, dtype=torch.int64, device=dev)
    for j in range(p):
        key = key * 4 + (reads[:, L - 1 - j].long() - 1)
    t0 = time.time()
    shards, sizes = [], []
    for s in range(args.sharded):
        idx = torch.nonzero(key == s).flatten()
        if idx.numel() == 0:
            continue
        b = dg.build_bwt(reads[idx], torch.full((idx.numel(),), L, dtype=torch.int32, device=dev))
        shards.append(engine.GpuBWT(runs=b.runs.cpu().numpy(), num_strings=b.num_strings, num_symbols=b.num_symbols,
                                    device=local_rank))
        sizes.append(int(shards[-1].info.hbm_bytes))
        del b, idx
    build_s = time.time() - t0
    Q, k = args.queries, args.k
    comm = sharding.connect_peers(engine.PeerComm(local_rank, 0, 1, Q), str(dev))

    def step():
        comm.begin(Q, stream)
        for S in shards:
            comm.accumulate(S, dq, k, Q, 0, stream)
        return comm.finish(stream)

    # parity: exact rank everywhere, shard sum == unsplit index
    for S in shards:
        S.set_tail_emulation(False)
    B_whole.set_tail_emulation(False)
    got = sharding.counts_view(step(), Q, dev).clone()
    want = torch.empty(Q, dtype=torch.int64, device=dev)
    B_whole.count_dev(dq, k, Q, 0, want, stream)
    torch.cuda.synchronize()
    ok = bool(torch.equal(got, want))
    for S in shards:
        S.set_tail_emulation(True)
    B_whole.set_tail_emulation(True)
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    ms = []
    for _ in range(5):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    t = sum(ms) / 1e3
    out = {"shards": len(shards), "split": f"reads' reversed last {p} bases (service.cpp:820)", "queries_per_step": Q,
           "population_queries_per_s": Q * len(ms) / t, "shard_queries_per_s": len(shards) * Q * len(ms) / t,
           "ms_per_step": 1e3 * t / len(ms), "kernels_per_step": len(shards) + 2,
           "shard_index_bytes_min_max": [min(sizes), max(sizes)], "sum_equals_unsplit_index_exact_rank": ok,
           "build_seconds": round(build_s, 1)}
    comm.close()
    for S in shards:
        S.close()
    assert ok, "sum of per-shard counts differs from the unsplit index"
    return out


def pinned_array(engine, nbytes, dtype, shape):
    import ctypes as C
    p = C.c_void_p()
    rc = engine.lib().rs_host_alloc(C.byref(p), nbytes)
    assert rc == 0
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    return np.frombuffer(buf, dtype=dtype).reshape(shape), p


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    from readserver_b200 import engine

    shard = build_shard(args, dev, rank)
    B = engine.GpuBWT(runs=shard["runs"], num_strings=shard["num_strings"], num_symbols=shard["num_symbols"],
                      device=local_rank)
    B.set_sai(shard["sai"])
    B.enable_stats(True)
    Q, k = args.queries, args.k
    dq = shard["queries"]
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()  # `ncu --profile-from-start off` skips the synthetic-data build
    dlo = torch.empty(Q, dtype=torch.int64, device=dev)
    dup = torch.empty(Q, dtype=torch.int64, device=dev)
    dcnt = torch.empty(Q, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream

    comm = None
    fused_ptr = [0]
    if world > 1:
        from readserver_b200 import sharding
        comm = sharding.connect_peers(engine.PeerComm(local_rank, rank, world, Q), dev)

    def step():
        if world > 1:
            # K1 adds every finished query's count into the owning rank's buffer over NVLink peer memory
            # (search + reduce-scatter in one kernel), then one flag barrier across the GPUs (C1)
            comm.begin(Q, stream)
            comm.accumulate(B, dq, k, Q, 0, stream)
            fused_ptr[0] = comm.finish(stream)
        else:
            B.find_interval_dev(dq, k, Q, dlo, dup, stream)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    B.read_stats(reset=True)
    if world > 1:
        dist.barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    torch.cuda.synchronize()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1)  # evict the index and the queries from L2 between timed steps (untimed)
        if world > 1:
            # the untimed flush ends at different moments on different GPUs: line the ranks up again so that
            # a step's time is the step, not the skew of what ran before it
            torch.cuda.synchronize()
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
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    interval_steps, _ = B.read_stats(reset=True)

    fused_ok = None
    if world > 1:
        # the fused reduce-scatter against NCCL: per-shard counts summed with an all-reduce, owned entries compared
        from readserver_b200 import sharding
        assert not comm.timed_out(), "peer barrier timed out"
        fused = sharding.counts_view(fused_ptr[0], Q, dev).clone()
        B.count_dev(dq, k, Q, 0, dcnt, stream)
        torch.cuda.synchronize()
        dist.all_reduce(dcnt)
        own = torch.from_numpy(sharding.owner_of(dq.cpu().numpy().reshape(Q, k), world)).to(dev) == rank
        ok = bool(torch.equal(fused[own], dcnt[own]) and int(fused[~own].abs().sum()) == 0)
        okt = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        fused_ok = bool(okt.item())
        assert fused_ok, "fused peer-memory reduction differs from the NCCL all-reduce of per-shard counts"

    # ---- e2e: the host-pointer C-ABI call, pinned host buffers, copies inside the timed region ----
    hq, _p1 = pinned_array(engine, Q * k, np.uint8, (Q, k))
    hlo, _p2 = pinned_array(engine, Q * 8, np.uint64, (Q,))
    hup, _p3 = pinned_array(engine, Q * 8, np.uint64, (Q,))
    hq[:] = dq.cpu().numpy()
    L = engine.lib()
    for _ in range(3):
        assert L.rs_find_interval_batch(B._h, hq.ctypes.data, k, Q, hlo.ctypes.data, hup.ctypes.data) == 0
    if world > 1:
        dist.barrier()
    e2e_s = 0.0
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = L.rs_find_interval_batch(B._h, hq.ctypes.data, k, Q, hlo.ctypes.data, hup.ctypes.data)
        e2e_s += time.perf_counter() - t0
        assert rc == 0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_t.item())

    # the same through count_reads' batch call (service.cpp:279-315): 8 bytes back per query instead of 16
    e2e_count = None
    if world == 1:
        hcnt, _p4 = pinned_array(engine, Q * 8, np.uint64, (Q,))
        for _ in range(2):
            assert L.rs_count_batch(B._h, hq.ctypes.data, k, Q, 0, hcnt.ctypes.data) == 0
        cs = 0.0
        nrep = max(3, args.steps // 4)
        for _ in range(nrep):
            flush.fill_(1)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rc = L.rs_count_batch(B._h, hq.ctypes.data, k, Q, 0, hcnt.ctypes.data)
            cs += time.perf_counter() - t0
            assert rc == 0
        e2e_count = Q * nrep / cs
        assert np.array_equal(hcnt, np.where(hup >= hlo, hup - hlo + 1, 0).astype(np.uint64)), "count != interval width"

    # results of the device path and of the host path agree (and, below, with the CPU reference)
    if world == 1:
        assert np.array_equal(dlo.cpu().numpy().view(np.uint64), hlo) and np.array_equal(dup.cpu().numpy().view(np.uint64), hup)

    # ---- side measurement: count + locate (findInterval -> rows -> extractPrefix walk -> .sai) ----
    locate = None
    if world == 1 and args.locate_queries > 0:
        QL = min(Q, args.locate_queries)
        max_hits, max_steps = 256, args.read_len + 28
        cap = QL * 128
        d_off = torch.empty(QL + 1, dtype=torch.int64, device=dev)
        d_ri = torch.empty(cap, dtype=torch.int64, device=dev)
        d_st = torch.empty(cap, dtype=torch.int32, device=dev)
        for _ in range(2):
            hits = B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri, d_st, stream)
        torch.cuda.synchronize()
        B.read_stats(reset=True)
        lms = []
        for _ in range(max(3, args.steps // 4)):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            hits = B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri, d_st, stream)
            e1.record()
            torch.cuda.synchronize()
            lms.append(e0.elapsed_time(e1))
        isteps, wsteps = B.read_stats(reset=True)
        lt = sum(lms) / 1e3
        locate = {"metric": "31-mer count+locate queries/s (interval, rows <= 256 per k-mer, LF walk to '$', .sai read index)",
                  "value": QL * len(lms) / lt, "unit": UNIT, "queries_per_step": QL, "hits_per_step": hits,
                  "hit_rows_per_s": hits * len(lms) / lt, "ms_per_step": 1e3 * lt / len(lms),
                  "walk_steps_per_step": wsteps / len(lms),
                  "achieved_GBps": (2.0 * isteps + wsteps) * BYTES_PER_RANK / lt / 1e9,
                  "kernels_per_step": 5, "note": "includes one stream sync per step to size the hit list"}
        # the same stage with the locate table (8 bytes per BWT row of HBM: one read per hit row instead of a walk)
        try:
            t0 = time.perf_counter()
            B.build_locate_table(args.read_len + 28)
            torch.cuda.synchronize()
            build_s = time.perf_counter() - t0
            d_ri2 = torch.empty(cap, dtype=torch.int64, device=dev)
            d_st2 = torch.empty(cap, dtype=torch.int32, device=dev)
            B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri2, d_st2, stream)
            torch.cuda.synchronize()
            same = bool(torch.equal(d_ri[:hits], d_ri2[:hits]) and torch.equal(d_st[:hits], d_st2[:hits]))
            tms = []
            for _ in range(max(3, args.steps // 4)):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri2, d_st2, stream)
                e1.record()
                torch.cuda.synchronize()
                tms.append(e0.elapsed_time(e1))
            tt = sum(tms) / 1e3
            locate["with_locate_table"] = {"value": QL * len(tms) / tt, "unit": UNIT, "hit_rows_per_s": hits * len(tms) / tt,
                                           "ms_per_step": 1e3 * tt / len(tms), "table_bytes": 8 * (int(shard["num_symbols"]) + int(shard["num_strings"])),
                                           "build_seconds": build_s, "equals_walk": same,
                                           "note": "rs_index_build_locate_table: where each row's walk ends, found once by "
                                                   "walking every read (n LF steps)"}
            # sampled variant: 1/16 of the rows hold an entry, a hit row is walked < 16 steps
            t0 = time.perf_counter()
            B.build_locate_table(args.read_len + 28, every=16)
            torch.cuda.synchronize()
            build16 = time.perf_counter() - t0
            tbytes16 = int(B.info.hbm_bytes)
            B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri2, d_st2, stream)
            torch.cuda.synchronize()
            same16 = bool(torch.equal(d_ri[:hits], d_ri2[:hits]) and torch.equal(d_st[:hits], d_st2[:hits]))
            sms = []
            for _ in range(max(3, args.steps // 4)):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                B.locate_dev(dq, k, QL, max_hits, max_steps, d_off, cap, None, None, d_ri2, d_st2, stream)
                e1.record()
                torch.cuda.synchronize()
                sms.append(e0.elapsed_time(e1))
            B.drop_locate_table()
            tbytes16 -= int(B.info.hbm_bytes)
            st_ = sum(sms) / 1e3
            locate["with_sampled_locate_table"] = {"every": 16, "value": QL * len(sms) / st_, "unit": UNIT,
                                                   "hit_rows_per_s": hits * len(sms) / st_, "ms_per_step": 1e3 * st_ / len(sms),
                                                   "table_bytes": tbytes16, "build_seconds": build16, "equals_walk": same16}
            del d_ri2, d_st2
        except engine.RsError as ex:
            locate["with_locate_table"] = {"unavailable": str(ex)}
        if shard.get("reads_text") is not None and hits:
            # output=reads from the packed read table, addressed by the read indices just located
            B.set_reads_arrays(shard["reads_text"], shard["reads_off"])
            ml = args.read_len + 4
            d_str = torch.empty(hits * ml, dtype=torch.uint8, device=dev)
            d_len = torch.empty(hits, dtype=torch.int32, device=dev)
            B.reads_dev(d_ri, hits, ml, d_str, d_len, stream)
            torch.cuda.synchronize()
            rms = []
            for _ in range(5):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                B.reads_dev(d_ri, hits, ml, d_str, d_len, stream)
                e1.record()
                torch.cuda.synchronize()
                rms.append(e0.elapsed_time(e1))
            # spot check against the table's source
            ri_h = d_ri[:64].cpu().numpy().astype(np.int64)
            got = d_str[: 64 * ml].cpu().numpy().reshape(64, ml)
            for t, r in enumerate(ri_h):
                if 0 <= r < shard["num_strings"]:
                    a, b = int(shard["reads_off"][r]), int(shard["reads_off"][r + 1])
                    assert np.array_equal(got[t, : b - a], shard["reads_text"][a:b]), "read table mismatch"
            locate["reads_from_table"] = {"reads_per_s": hits * len(rms) / (sum(rms) / 1e3), "ms": sum(rms) / len(rms),
                                          "bytes_written": hits * ml,
                                          "note": "rs_reads_batch_dev on the located read indices (replaces extractPrefix + "
                                                  "extractPostfix per row, query.cpp:87-100)"}
            del d_str, d_len
        if hits:
            # config-5 style tail: hit reads -> sample codes -> sorted distinct sample set per k-mer (K3), through the
            # host-pointer call (copies inside); synthetic table: 100 samples, 1-3 codes per read
            rng = np.random.default_rng(11)
            ns = int(shard["num_strings"])
            soff = np.zeros(ns + 1, dtype=np.uint64)
            np.cumsum(rng.integers(1, 4, ns), out=soff[1:])
            scodes = (rng.integers(0, 100, int(soff[-1])) + 34).astype(np.uint8)   # code of sample s: s + 1 + 33
            B.set_samples(soff, scodes, 1)
            h_off = d_off.cpu().numpy().view(np.uint64)
            h_ri = d_ri[:hits].cpu().numpy().view(np.uint64)
            B.carriers(h_off, h_ri)
            cs = []
            for _ in range(3):
                t0 = time.perf_counter()
                coff, ccodes = B.carriers(h_off, h_ri)
                cs.append(time.perf_counter() - t0)
            locate["sample_sets"] = {"kmers_per_s": QL / min(cs), "hit_reads_per_s": hits / min(cs), "ms": 1e3 * min(cs),
                                     "samples_out": int(coff[-1]), "note": "rs_carriers_batch (host pointers, two passes: "
                                     "sizes, then codes): RocksDB Get + decode + set union per k-mer in the reference "
                                     "(service.cpp:816-845)"}
            del soff, scodes
        del d_off, d_ri, d_st

    sharded = None
    if world == 1 and args.sharded > 0 and shard.get("reads_dev") is not None:
        sharded = sharded_measure(args, shard, B, dq, engine, local_rank, stream, flush)
        shard["reads_dev"] = None
        torch.cuda.empty_cache()

    random_gbs = random_gbs_128 = None
    if world == 1 and not args.no_random_roofline:
        try:
            random_gbs = engine.measure_random_access(local_rank, 8 << 30, 1 << 27, 64)
            random_gbs_128 = engine.measure_random_access(local_rank, 8 << 30, 1 << 27, 128)
        except Exception as ex:  # not fatal for the bench line
            random_gbs = None
            print(f"random-access roofline measurement failed: {ex}", file=sys.stderr)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = world * Q * args.steps / (total_ms / 1e3)
    peak, peak_src = hbm_peak()
    traffic, traffic_src = k1_traffic()
    kern_ms = sum(step_ms) / len(step_ms)  # at N = 1 the step IS the kernel launch
    ranks_per_launch = 2.0 * interval_steps / args.steps
    achieved = ranks_per_launch * BYTES_PER_RANK / (kern_ms / 1e3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {
            "workload": workload_name(args), "k": k, "queries_per_step": Q, "shards_per_gpu": 1,
            "threads_per_query": 1, "block_layout": "64 B = 28 B header (start, 4 x 40-bit counters) + 36 run bytes; 2 sectors per rank",
            "bwt_symbols": shard["num_symbols"], "bwt_runs": int(shard["runs"].size), "reads_unique": shard["num_strings"],
            "index_hbm_bytes": int(B.info.hbm_bytes), "index_fits_l2": bool(B.info.hbm_bytes < 126e6),
            "l2": "256 MiB buffer written between timed steps (untimed) -> index and queries start cold",
            "timing": "CUDA events around each step on the launching stream, summed; max over ranks"
                      + ("; ranks re-aligned with a barrier after each untimed L2 flush" if world > 1 else ""),
            "collective": "none" if world == 1 else
            "fused into K1: counts added into the owning rank's buffer over NVLink peer memory + one flag barrier "
            "per step, inside the timed step; checked against an NCCL all-reduce of the per-shard counts",
            "fused_reduce_equals_nccl": fused_ok,
            "data_build_seconds": round(shard["build_seconds"], 1),
        },
        "roofline": {
            "bound": "hbm", "kernel": "k_find_interval", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "algorithmic_bytes": "executed getOcc calls x 64 B (one 64-byte block: header + 36 run bytes); "
                                 "lower/upper of a step often share a block",
            "getocc_per_launch": ranks_per_launch,
            "achieved_at_128B_per_getocc": 2 * achieved, "frac_at_128B_per_getocc": 2 * achieved / peak,
            "random_access_peak": random_gbs,
            "frac_of_random_access_peak": (achieved / random_gbs) if random_gbs else None,
            "random_access_peak_source": "measured in this run: random aligned 64-B loads (one block each) over 8 GiB "
                                         "(rs_measure_random_access_granule)",
            "random_access_peak_128B_lines": random_gbs_128,
        },
        "e2e": {"value": world * Q * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": Q * k,
                "d2h_bytes_per_step": 2 * Q * 8, "api": "rs_find_interval_batch (host pointers, pinned)",
                "count_api_value": e2e_count,
                "count_api": "rs_count_batch, forward strand: 8 bytes back per query (not the headline: kept beside it)"},
        "gpu_launches": args.steps * 3,  # k_pack_kmers + k_find_interval + k_search_flagged (N > 1: k_peer_barrier)
        "clocks": clk,
    }
    if locate:
        line["locate"] = locate
    if sharded:
        line["sharded"] = sharded
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        r = cpu_find_interval(shard, hq, k, cores, target_seconds=args.cpu_seconds, steps=1, extras=True)
        ok = bool(np.array_equal(r["lower"], hlo[: r["n"]]) and np.array_equal(r["upper"], hup[: r["n"]]))
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": cores, "kind": r["kind"],
                                "sample": f"first {r['n']} of {Q} queries, {cores} threads striding the array",
                                "gpu_results_bit_exact_on_sample": ok,
                                "as_shipped_O0_value": r.get("as_shipped_O0_value"),
                                "locate_rows_per_s": r.get("locate_rows_per_s"), "locate_rows": r.get("locate_rows"),
                                "note": "value: reference sources at -O2; as_shipped_O0_value: built without -O as its "
                                        "Makefiles do; locate_rows_per_s: extractPrefix per hit row on the same threads"}
        assert ok, "GPU results differ from the CPU reference"
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = sys.stdout


def main():
    args = parse()
    # libraries (NCCL's version banner, torchrun notices) must not share stdout with the ONE JSON line:
    # everything but that line goes to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
