#!/bin/bash
# Rebuild the library with different K1 occupancy targets on the GPU box and time the bench's device leg.
O=gpurun_out; mkdir -p $O
for b in "$@"; do
  RS_NVCC_EXTRA="-DRS_K1_BLOCKS=$b" python -m readserver_b200.build --force > /dev/null 2>&1 || { echo "build $b failed"; continue; }
  if [ "$b" == "$1" ]; then timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2; fi
  timeout 300 python bench.py --no-cpu-baseline --no-random-roofline --locate-queries 0 > $O/tune_$b.json 2> $O/tune_$b.err
  python - <<PY
import json
d=json.load(open("$O/tune_$b.json"))
print("K1_BLOCKS=$b value %.1f M q/s  e2e %.1f M  ms %.4f" % (d["value"]/1e6, d["e2e"]["value"]/1e6, d["ms_per_step"]))
PY
done
