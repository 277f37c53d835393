"""Golden vectors from the reference's own code for the rows that round 1 left unpinned (run in the build container,
where /root/reference exists; `make -C oracle ref` first):

  gt_golden.json       find_gt_reads (src/service/service.cpp:507-711, compiled from the reference's source through
                       oracle/ref_service_shim.cpp) per tile of a set of site windows over tests/golden/clean.bwt
  sample_codes.json    src/util/create_samples_hash.cpp run on name lists of several sizes
  config_golden.json   a worker .cfg (this repository's own text, the syntax of demo/TEMPLATE.service.cfg) as libconfig 1.5
                       reads it

    python -m tests.golden.make_golden2
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import gt as ogt  # noqa: E402
from tests import common  # noqa: E402

GT_GRID = [(31, 7, 10000), (31, 7, 12), (31, 0, 3), (12, 5, 40), (20, 19, 1), (60, 9, 6)]

CFG_TEXT = '''// worker configuration in the syntax of the reference's template
# a hash comment
/* a block
   comment */
prefix = "/data/bwt/final"
hashfile = "/data/files/list_of_sample_hash";
pull = "tcp://10.0.0.1:5555";
push = "tcp://10.0.0.1:5556";
push_count = "tcp://10.0.0.1:5557";
suffix = "";
size_of_sample = "1";
has_other_meta_data = "0";
max_read_length = "100";
min_read_length = "73"
rocksdb_path = "/data/rocksdbs/";   // trailing comment
rocksdb_ext = ".rocksdb";
escaped = "a \\"quoted\\" word\\tand a tab";
adjacent = "two " "parts";
rocksdb = [
    "CGA",
    "CGC", "CGG" ,
    "TTT"
];
'''
CFG_STRINGS = ["prefix", "hashfile", "pull", "push", "push_count", "suffix", "size_of_sample", "has_other_meta_data",
               "max_read_length", "min_read_length", "rocksdb_path", "rocksdb_ext", "escaped", "adjacent", "absent"]
CFG_ARRAYS = ["rocksdb", "absent_list"]


def gt_windows(case, kmers, max_iv):
    from tests.test_gt import windows_of
    wins = windows_of(case, 24, seed=kmers * 1000 + max_iv)
    wins += [(wins[1][0][:kmers], kmers // 2 + 1), (wins[2][0][:kmers + 3], 2), (wins[6][0][-kmers - 2:], kmers + 2)]
    return wins


def digest(strings):
    return hashlib.sha1("\n".join(strings).encode()).hexdigest()


def main():
    oracle.build(ref=True)
    assert oracle.have_ref_service(), "needs /root/reference"
    ds = common.Datasets()
    case, P = ds.case("clean"), ds.port("clean")
    R = oracle.RefService(os.path.join(HERE, "clean.bwt"))
    gt = []
    for kmers, skip, max_iv in GT_GRID:
        R.set_limits(max_iv)
        for w, pos in gt_windows(case, kmers, max_iv):
            tiles = (len(w) - kmers) // (skip + 1) + 1 if len(w) >= kmers and pos <= len(w) else 0
            per = []
            for i in range(tiles):
                try:   # the restatement only decides which tiles the reference never returns from
                    ogt.find_gt_reads(P, w, i, kmers, skip, pos, max_iv)
                except ogt.GtUndefined as e:
                    if e.kind == "hang":
                        per.append("hang")
                        continue
                ref = R.find_gt_reads(w, i, kmers, skip, pos)
                per.append("throw" if ref is None else [len(ref), digest(ref)])
            gt.append({"kmers": kmers, "skip": skip, "max_interval_size": max_iv, "w": w, "pos": pos, "tiles": per})
    json.dump({"bwt": "clean.bwt", "windows": gt}, open(os.path.join(HERE, "gt_golden.json"), "w"))
    print("gt_golden:", len(gt), "windows,", sum(len(g["tiles"]) for g in gt), "tiles")

    codes = {}
    for n in (5, 222, 223, 1000, 60000):
        names = [f"sample_{i:06d}" for i in range(n)]
        m = oracle.ref_create_samples_hash(names)
        ordered = [m[x] for x in sorted(names)]
        width = len(ordered[0])
        assert all(len(c) == width for c in ordered)
        blob = b"".join(ordered)
        codes[str(n)] = {"width": width, "sha256": hashlib.sha256(blob).hexdigest(),
                         "first": [c.hex() for c in ordered[:4]], "last": [c.hex() for c in ordered[-4:]],
                         "all_hex": blob.hex() if n <= 1000 else None}
    json.dump(codes, open(os.path.join(HERE, "sample_codes.json"), "w"))
    print("sample_codes:", {k: v["width"] for k, v in codes.items()})

    cfg_path = os.path.join(HERE, "worker.cfg")
    with open(cfg_path, "w") as f:
        f.write(CFG_TEXT)
    parsed = oracle.ref_config(cfg_path, CFG_STRINGS, CFG_ARRAYS)
    json.dump(parsed, open(os.path.join(HERE, "config_golden.json"), "w"), indent=1)
    print("config_golden:", parsed)


if __name__ == "__main__":
    main()
