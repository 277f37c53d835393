"""Shards across the GPUs of one box.

The reference runs one worker process per BWT shard; the front end broadcasts every query to every
worker (ZMQ PUB, ``src/service/server.cpp:124,957``) and sums the per-shard counts / concatenates the
per-shard read lists as replies arrive (``server.cpp:184-197, 404-410``).  Here the shards of a box are
spread over its GPUs (one process per GPU), every rank searches the whole query batch in the shards
it holds, and the fan-in becomes collectives over NVLink:

  C1  per-shard counts  -> sum              one all-reduce (or reduce-scatter to the k-mer's owner)
  C2  per-shard hit lists -> concatenation  size exchange + all-gather of the padded lists

``torch.distributed`` is the plumbing (NCCL on GPUs; the same code runs on gloo for the CPU tests).
Which rank owns the reduced result of a k-mer is decided by the k-mer's prefix: with read-partitioned
shards that is the only place where routing by prefix is meaningful (a k-mer can occur in any shard).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def assign_shards(shard_bytes: list[int], world: int) -> list[list[int]]:
    """Longest-processing-time assignment of shards to ranks, balanced by bytes in HBM.
    Returns, per rank, the list of shard numbers it holds (deterministic)."""
    order = sorted(range(len(shard_bytes)), key=lambda i: (-shard_bytes[i], i))
    load = [0] * world
    out: list[list[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda j: (load[j], j))
        out[r].append(i)
        load[r] += shard_bytes[i]
    for lst in out:
        lst.sort()
    return out


def permutation_keys(k: int = 3) -> list[str]:
    """The shard keys of demo/permutations-3.txt: all strings of length k over ACGT, in that file's order."""
    keys = [""]
    for _ in range(k):
        keys = [p + c for p in keys for c in "ACGT"]
    return keys


def shard_of_read(read: str, k: int = 3) -> str:
    """A read's shard key: its last k bases, reversed (``string(it->rbegin(), it->rbegin()+3)``,
    src/service/service.cpp:820; src/util/load_data_into_rocksdb.cpp:45)."""
    return read[::-1][:k]


def owner_of(kmers: np.ndarray, world: int, prefix_len: int = 4) -> np.ndarray:
    """Rank that owns the reduced result of each k-mer ([Q, k] ASCII): its first prefix_len bases read
    as a base-4 number (any byte outside ACGT counts as A), modulo the number of ranks."""
    lut = np.zeros(256, dtype=np.int64)
    for i, c in enumerate(b"ACGT"):
        lut[c] = i
    p = min(prefix_len, kmers.shape[1])
    v = np.zeros(kmers.shape[0], dtype=np.int64)
    for j in range(p):
        v = v * 4 + lut[kmers[:, j]]
    return (v % world).astype(np.int32)


def reduce_counts(local_counts: torch.Tensor, group=None) -> torch.Tensor:
    """C1: element-wise sum of the ranks' per-query counts (int64 [Q]); every rank gets the totals."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(local_counts, op=dist.ReduceOp.SUM, group=group)
    return local_counts


def reduce_scatter_counts(local_counts: torch.Tensor, owners: torch.Tensor, group=None) -> torch.Tensor:
    """C1 with prefix-routed ownership: returns the totals of the queries this rank owns (in query
    order).  owners[q] = owning rank (owner_of).  Queries are permuted so that each owner's block is
    contiguous and equally padded, reduce-scattered, and un-padded."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local_counts
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = [torch.nonzero(owners == r).flatten() for r in range(world)]
    width = max(int(p.numel()) for p in per)
    stacked = torch.zeros((world, width), dtype=local_counts.dtype, device=local_counts.device)
    for r, idx in enumerate(per):
        stacked[r, : idx.numel()] = local_counts[idx]
    out = torch.empty(width, dtype=local_counts.dtype, device=local_counts.device)
    if dist.get_backend(group) == "gloo":  # gloo has no reduce_scatter: all-reduce and take our row
        dist.all_reduce(stacked, group=group)
        out = stacked[rank].clone()
    else:
        dist.reduce_scatter_tensor(out, stacked.reshape(-1), op=dist.ReduceOp.SUM, group=group)
    return out[: per[rank].numel()]


def gather_hits(hit_offsets: torch.Tensor, hits: torch.Tensor, group=None):
    """C2: concatenate every rank's per-query hit lists (CSR: hit_offsets [Q+1], hits [H]) so that each
    rank holds, per query, rank 0's hits, then rank 1's, ... (the order in which the reference's front
    end would see shard replies is arbitrary; this one is deterministic).
    Returns (offsets [Q+1], hits, src_rank [H])."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return hit_offsets, hits, torch.zeros_like(hits, dtype=torch.int32)
    world = dist.get_world_size(group)
    dev = hits.device
    q = hit_offsets.numel() - 1
    counts = (hit_offsets[1:] - hit_offsets[:-1]).contiguous()
    all_counts = [torch.empty_like(counts) for _ in range(world)]
    dist.all_gather(all_counts, counts, group=group)
    sizes = torch.tensor([int(c.sum()) for c in all_counts], device=dev)
    width = int(sizes.max())
    padded = torch.zeros(max(width, 1), dtype=hits.dtype, device=dev)
    padded[: hits.numel()] = hits
    all_hits = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(all_hits, padded, group=group)
    cnt = torch.stack(all_counts)                                 # [world, Q]
    tot = cnt.sum(0)
    out_off = torch.zeros(q + 1, dtype=hit_offsets.dtype, device=dev)
    torch.cumsum(tot, 0, out=out_off[1:])
    out = torch.empty(int(out_off[-1]), dtype=hits.dtype, device=dev)
    src = torch.empty(int(out_off[-1]), dtype=torch.int32, device=dev)
    within = torch.cumsum(cnt, 0) - cnt                           # start of rank r's hits inside query q
    for r in range(world):
        c = cnt[r]
        n = int(c.sum())
        if not n:
            continue
        qid = torch.repeat_interleave(torch.arange(q, device=dev), c, output_size=n)
        src_start = torch.cumsum(c, 0) - c
        j = torch.arange(n, device=dev) - src_start[qid]
        dst = out_off[:-1][qid] + within[r][qid] + j
        out[dst] = all_hits[r][:n]
        src[dst] = r
    return out_off, out, src


class ShardedCounter:
    """The shards one rank holds + the collective that turns per-shard counts into population counts.

    `count_fns` are callables kmers[Q,k] (uint8 numpy) -> int64 counts [Q], one per local shard; on a GPU
    rank they are ``GpuBWT.count`` bound methods (or the *_dev form writing into torch buffers).
    """

    def __init__(self, count_fns, device="cpu", group=None):
        self.count_fns = list(count_fns)
        self.device = device
        self.group = group

    def count(self, kmers: np.ndarray, strand: int = 0) -> torch.Tensor:
        total = torch.zeros(kmers.shape[0], dtype=torch.int64, device=self.device)
        for f in self.count_fns:
            c = f(kmers, strand)
            c = torch.as_tensor(np.asarray(c).astype(np.int64)) if not torch.is_tensor(c) else c
            total += c.to(self.device)
        return reduce_counts(total, self.group)


def connect_peers(comm, device, group=None):
    """Exchange the CUDA IPC handles of a `engine.PeerComm` over torch.distributed and map the peers."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    mine = torch.from_numpy(comm.export()).to(device)
    if world > 1:
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine, group=group)
        handles = torch.cat(allh).cpu().numpy()
    else:
        handles = mine.cpu().numpy()
    comm.connect(handles)
    if world > 1:
        dist.barrier(group=group)  # nobody searches before every rank has mapped every peer
    return comm


def counts_view(ptr: int, q: int, device) -> torch.Tensor:
    """int64 torch view of the Q counts at device address `ptr` (the buffer belongs to the communicator)."""
    class _Arr:
        __cuda_array_interface__ = {"shape": (q,), "typestr": "<i8", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_Arr(), device=device)
