"""rs_count_batch_noalloc_dev: the count path with its scratch in the library's static device array (no cudaMalloc).

Collected LAST on purpose: the entry point has no counterpart in the reference;
it is checked against the existing entry point rs_count_batch_dev,
bit for bit (integer work), on calls in a row that share the static array."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from readserver_b200 import engine as e
    return e


def _queries(case, n, k, seed):
    q = np.ascontiguousarray(case.queries(n, k, seed=seed)).reshape(n, k).copy()
    rng = np.random.default_rng(seed)
    bad = rng.choice(n, size=max(1, n // 200), replace=False)  # bytes outside ACGT: count 0 (service.cpp:299-301)
    q[bad, rng.integers(0, k, size=bad.size)] = ord("N")
    return q


@pytest.mark.parametrize("dense,table", [(True, 0), (True, 8), (False, 0), (False, 8)])
def test_noalloc_count_equals_count(datasets, engine, dense, table):
    import torch
    case = datasets.case("mid")
    dev = torch.device("cuda:0")
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        B.set_dense(dense)
        B.set_prefix_table(table)
        # everything the calls touch exists before the first of them
        calls = [(60000, 31, 0, 7), (20000, 31, 1, 8), (45000, 12, 0, 9), (300, 1, 1, 10)]  # (Q, k, strand, seed)
        bufs = []
        for Q, k, strand, seed in calls:
            d_q = torch.from_numpy(_queries(case, Q, k, seed)).to(dev)
            want = torch.full((Q,), 0x5A5A5A5A, dtype=torch.int64, device=dev)
            got = torch.full((Q,), 0x3C3C3C3C, dtype=torch.int64, device=dev)
            bufs.append((d_q, want, got))
        for (Q, k, strand, _), (d_q, want, got) in zip(calls, bufs):
            B.count_dev(d_q, k, Q, strand, want)  # the existing entry point
        torch.cuda.synchronize()
        # two (and more) calls in a row on one stream: each finds the previous call's keys and counter in the array;
        # the second is smaller than the first, so stale keys lie beyond its Q
        free_before = None
        for n, ((Q, k, strand, _), (d_q, want, got)) in enumerate(zip(calls, bufs)):
            B.count_noalloc_dev(d_q, k, Q, strand, got)
            if n == 0:  # the first call took the symbol's address; from here on device memory must not move
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info(0)[0]
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(0)[0] >= free_before, "a no-alloc call took device memory"
        for (Q, k, strand, _), (d_q, want, got) in zip(calls, bufs):
            assert torch.equal(got, want), (Q, k, strand)
        assert int(bufs[0][1].sum()) > 0  # the batch does hit
        # and the same call again, over its own leftovers
        d_q, want, got = bufs[0]
        got.fill_(-1)
        B.count_noalloc_dev(d_q, 31, calls[0][0], 0, got)
        torch.cuda.synchronize()
        assert torch.equal(got, want)


def test_noalloc_limits(datasets, engine):
    import torch
    case = datasets.case("small")
    dev = torch.device("cuda:0")
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        Q = engine.NOALLOC_MAX_QUERIES + 1
        d_q = torch.full((Q * 4,), ord("A"), dtype=torch.uint8, device=dev)
        out = torch.zeros(Q, dtype=torch.int64, device=dev)
        with pytest.raises(engine.RsError) as ei:
            B.count_noalloc_dev(d_q, 4, Q, 0, out)
        assert "exceed" in str(ei.value) and str(engine.NOALLOC_MAX_QUERIES) in str(ei.value)
        with pytest.raises(engine.RsError):
            B.count_noalloc_dev(d_q, 32, 10, 0, out)   # longer than a packed key
        with pytest.raises(engine.RsError):
            B.count_noalloc_dev(d_q, 4, 10, 2, out)    # strand
        assert int(out.sum()) == 0                      # nothing was launched
        # the maximum itself is accepted and agrees
        Q = engine.NOALLOC_MAX_QUERIES
        q = torch.from_numpy(np.ascontiguousarray(case.queries(Q, 4, seed=3))).to(dev)
        want = torch.empty(Q, dtype=torch.int64, device=dev)
        got = torch.empty(Q, dtype=torch.int64, device=dev)
        B.count_dev(q, 4, Q, 0, want)
        B.count_noalloc_dev(q, 4, Q, 0, got)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
