"""One rank of the multi-process shard-set test (tests/test_gpu_pop.py::test_multi_rank_fan_in_through_peer_memory;
also runnable under torchrun on a multi-GPU box).  TEST INFRASTRUCTURE.

Every rank builds the same small population deterministically, keeps the shards placed on it, joins the ranks'
communicators through CUDA IPC (handles exchanged over gloo) and runs the same batch; then checks the answers of the
k-mers it owns against the per-shard CPU oracles."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from readserver_b200 import engine, sharding
    from tests.test_gpu_pop import PopCase
    locate_every = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    case = PopCase(engine, device=dev)
    q = case.queries[:2600]
    Q = q.shape[0]
    placement = sharding.assign_shards([int(x) for x in case.hist.cpu().numpy()], world)
    cap = 1 << 18
    S = engine.ShardSet(dev)
    for s in placement[rank]:
        S.add(case.gpu[s], s)
    S.seal(case.read_len + 28, locate_every)
    comm = sharding.connect_peers(engine.PopComm(dev, rank, world, Q, cap), f"cuda:{dev}")
    dq = torch.from_numpy(q).to(f"cuda:{dev}")
    carriers = locate_every == 1
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):                     # twice: the buffers are cleared and reused correctly
        comm.query_dev(S, dq, 31, Q, False, 1 << 20, True, carriers, stream)
        torch.cuda.synchronize()
        dist.barrier()
        assert not comm.timed_out(), "a peer barrier timed out"
    order = [s for r in range(world) for s in placement[r]]
    want = case.expected(q, 1 << 20, shard_order=order)
    owner = sharding.owner_of(q, world)
    res = {k: v.cpu().numpy().view(np.uint64 if k != "carriers" else np.uint32) for k, v in comm.results(Q).items()}
    mine = owner == rank
    assert np.array_equal(res["counts"][mine], want[0][mine]) and int(res["counts"][~mine].sum()) == 0
    off = res["hit_offsets"]
    assert int(res["total"][0]) == int(off[Q]) <= cap
    seen = 0
    for i in range(Q):
        a, b = int(off[i]), int(off[i + 1])
        if not mine[i]:
            assert a == b
            continue
        wa, wb = int(want[1][i]), int(want[1][i + 1])
        assert b - a == wb - wa and np.array_equal(res["hits"][a:b], want[2][wa:wb]), (rank, i)
        seen += b - a
    if carriers:
        assert np.array_equal(res["carriers"][mine], want[3][mine]) and int(res["carriers"][~mine].sum()) == 0
    tot = torch.tensor([seen])
    dist.all_reduce(tot)
    assert int(tot) == int(want[1][Q]) > Q
    # count-only call: the fused count reduction alone
    comm.query_dev(S, dq, 31, Q, False, 1 << 20, False, False, stream)
    torch.cuda.synchronize()
    dist.barrier()
    c2 = comm.results(Q)["counts"].cpu().numpy().view(np.uint64)
    assert np.array_equal(c2[mine], want[0][mine])
    dist.barrier()
    comm.close()
    S.close()
    case.close()
    dist.destroy_process_group()
    print("POP_WORKER_OK", rank, flush=True)


if __name__ == "__main__":
    main()
