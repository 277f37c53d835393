"""Seed stage of /gt (find_gt_reads, service.cpp:507-711): the per-tile seed logic the kernel runs (rs_gt_ops.h,
replayed on the host by tests/emul) and, on a GPU, the whole batch call, against the CPU restatement oracle/gt.py
built on the oracle's findInterval / extractPrefix / extractPostfix."""
import numpy as np
import pytest

from oracle import gt as ogt
from readserver_b200 import datagen as dg
from tests import common

KIND = {"all": 0, "postfix": 1, "prefix": 2}
FLAG = {"throw": 1, "hang": 2}


def windows_of(case, n, seed, max_read=100):
    """(window, pos) pairs cut the way GtTask::run cuts them (service.cpp:1027-1037): up to max_read-1 bases
    either side of the site, site position 1-based inside the window."""
    g = bytes(dg.codes_to_ascii(case.genome)).decode()
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        site = int(rng.integers(1, len(g) + 1))          # 1-based in the genome
        if i % 5 == 0:
            site = int(rng.integers(1, max_read))         # near the start: short left flank
        if site <= max_read:
            w, pos = g[:max_read + site - 1], site
        else:
            w, pos = g[site - max_read:site - max_read + 2 * max_read - 1], max_read
        if i % 7 == 3:
            w = w[:len(w) // 2 + 3]                        # short right flank
            pos = min(pos, len(w))
        if i % 11 == 4:
            j = int(rng.integers(0, len(w)))
            w = w[:j] + "N" + w[j + 1:]
        out.append((w, pos))
    return out


@pytest.mark.parametrize("kmers,skip,max_iv", [(31, 7, 10000), (31, 7, 12), (31, 0, 3), (12, 5, 40), (20, 19, 1), (60, 9, 6)])
def test_tile_seeds_match_restated_reference(kmers, skip, max_iv, datasets):
    case, P = datasets.case("clean"), datasets.port("clean")
    E = common.Emul(case)
    seen = {"wide": 0, "throw": 0, "hang": 0, "two": 0, 0: 0, 1: 0, 2: 0}
    wins = windows_of(case, 24, seed=kmers * 1000 + max_iv)
    # one-tile windows: an extension runs into both ends (the reference hangs in its first loop and never gets
    # to the second one, which would throw)
    wins += [(wins[1][0][:kmers], kmers // 2 + 1), (wins[2][0][:kmers + 3], 2), (wins[6][0][-kmers - 2:], kmers + 2)]
    for w, pos in wins:
        got = E.gt_seeds(w, pos, kmers, skip, max_iv)
        tiles = (len(w) - kmers) // (skip + 1) + 1 if len(w) >= kmers and pos <= len(w) else 0
        assert got.shape[0] == tiles
        for i in range(tiles):
            lo0, hi0, thr0, lo1, hi1, thr1, kind, flag = (int(x) for x in got[i])
            seeds = []
            try:
                ogt.find_gt_reads(P, w, i, kmers, skip, pos, max_iv, seeds=seeds)
            except ogt.GtUndefined as e:
                assert flag == FLAG[e.kind], (w, pos, i)
                assert lo0 > hi0 and lo1 > hi1
                seen[e.kind] += 1
                continue
            assert flag == 0, (w, pos, i)
            mine = [(lo0, hi0, thr0), (lo1, hi1, thr1)]
            assert len(seeds) <= 2
            for j in range(2):
                if j >= len(seeds):
                    assert mine[j][0] > mine[j][1]
                    continue
                lo, hi, k, thr = seeds[j]
                assert (mine[j][0], mine[j][1]) == (lo, hi), (w, pos, i, j)
                if hi >= lo:
                    assert kind == KIND[k]
                    if k != "all":
                        assert mine[j][2] == thr
                    seen[kind] += 1
            seen["two"] += len(seeds) == 2
    # every branch of the reference function was exercised somewhere in the parameter grid
    if max_iv == 10000:
        assert seen[0] and seen[1] and seen[2] and not seen["two"]
    if max_iv <= 12:
        assert seen["two"] > 0


def test_restatement_agrees_with_brute_force(datasets):
    """Without extensions (intervals below max_interval_size) the read set of a tile is: reads holding the tile
    that reach within 4 bases of the site."""
    case, P = datasets.case("clean"), datasets.port("clean")
    kmers, skip = 31, 7
    for w, pos in windows_of(case, 6, seed=77):
        if "N" in w:
            continue
        tiles = (len(w) - kmers) // (skip + 1) + 1
        for i in range(tiles):
            start = (skip + 1) * i + 1
            end = start + kmers - 1
            tile = w[start - 1:end]
            want = set()
            for r, read in enumerate(case.reads):
                o = read.find(tile)
                while o >= 0:
                    after = len(read) - o - kmers        # bases following the tile in the read
                    if pos > end:
                        keep = not (pos - end > after + 4)
                    elif pos < start:
                        keep = not (start - pos > o + 4)
                    else:
                        keep = True
                    if keep:
                        want.add(r)
                    o = read.find(tile, o + 1)
            assert ogt.find_gt_reads(P, w, i, kmers, skip, pos) == want


@pytest.mark.gpu
@pytest.mark.parametrize("kmers,skip,max_iv", [(31, 7, 10000), (31, 7, 12), (31, 0, 3), (12, 5, 40), (60, 9, 6)])
def test_gt_reads_batch_on_gpu(kmers, skip, max_iv, datasets):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from readserver_b200 import engine
    case, P = datasets.case("clean"), datasets.port("clean")
    wins = windows_of(case, 40, seed=kmers * 1000 + max_iv)
    wins += [("ACGT", 2), ("", 0), (wins[0][0], len(wins[0][0]) + 1)]          # shorter than a tile; empty; pos > size
    wins += [(wins[1][0][:kmers], kmers // 2 + 1), (wins[2][0][:kmers + 3], 2), (wins[6][0][-kmers - 2:], kmers + 2)]
    want = [ogt.gt_task_reads(P, w, pos, kmers, skip, max_iv) for w, pos in wins]
    with engine.GpuBWT(runs=case.runs, num_strings=case.num_strings, num_symbols=case.num_symbols) as B:
        with pytest.raises(engine.RsError):
            B.gt_reads([w for w, _ in wins], [p for _, p in wins], kmers, skip, max_iv)     # no .sai yet
        B.set_sai(case.sai)
        for table in (False, True):
            if table:
                B.set_reads(case.reads)       # postfix lengths from the packed read table instead of the select walk
            off, ri, flags, cand = B.gt_reads([w for w, _ in wins], [p for _, p in wins], kmers, skip, max_iv)
            assert cand >= ri.size
            for q, (reads, undefined) in enumerate(want):
                got = [int(x) for x in ri[int(off[q]):int(off[q + 1])]]
                assert got == reads, (q, wins[q], table)
                assert int(flags[q]) == sum(FLAG[k] for k in undefined), (q, wins[q])
        with pytest.raises(engine.RsError) as e:
            B.gt_reads([w for w, _ in wins], [p for _, p in wins], kmers, skip, max_iv, max_rows=1)
        assert e.value.status == 9
    assert sum(len(r) for r, _ in want) > 50


@pytest.mark.parametrize("name", ["tiny", "small", "clean", "mid"])
def test_stray_bytes_searched_as_the_reference_does(name, datasets):
    """A byte outside "$ACGT" takes the '$' counters of the reference's tree but matches no run in its final scan
    (rlebwt.cpp:286,293): getOcc = '$' before the 64-run bucket.  The device code recovers that boundary from its
    own blocks (rs_occ_stray); checked here through findInterval on strings with such bytes at every place."""
    case, P = datasets.case(name), datasets.port(name)
    E = common.Emul(case)
    rng = np.random.default_rng(5)
    for k in (1, 2, 5, 31, 60):
        q = case.queries(300, k, seed=40 + k).copy()
        for i in range(q.shape[0]):
            for _ in range(1 + i % 3):
                q[i, int(rng.integers(0, k))] = [ord("N"), ord("$"), ord("n"), 0, 255, ord("X")][i % 6]
        want = P.find_interval_batch(q, k)
        got = E.find_interval(q, k)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (name, k)
