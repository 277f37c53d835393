#!/bin/bash
# scripts/tune_macro.sh MACRO v1 v2 ...: rebuild with -DMACRO=v on the GPU box and print the bench's device numbers.
O=gpurun_out; mkdir -p $O; M=$1; shift
for b in "$@"; do
  RS_NVCC_EXTRA="-D$M=$b" python -m readserver_b200.build --force > /dev/null 2>&1 || { echo "build $b failed"; continue; }
  timeout 300 python bench.py --steps 20 --no-cpu-baseline --no-random-roofline > $O/tune_${M}_$b.json 2> $O/tune_${M}_$b.err
  python - <<PY
import json
d=json.loads(open("$O/tune_${M}_$b.json").read().strip().splitlines()[-1])
print("$M=$b value %.1f M q/s  e2e %.1f M  locate rows %.1f M/s" % (d["value"]/1e6, d["e2e"]["value"]/1e6, d["locate"]["hit_rows_per_s"]/1e6))
PY
done
