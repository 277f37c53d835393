#!/bin/bash
# scripts/tune_flags.sh "<nvcc -D flags>" ...: rebuild with each flag set on the GPU box, print the bench's device numbers.
O=gpurun_out; mkdir -p $O; i=0
for fl in "$@"; do
  i=$((i+1))
  RS_NVCC_EXTRA="$fl" python -m readserver_b200.build --force > /dev/null 2>&1 || { echo "build [$fl] failed"; continue; }
  timeout 300 python bench.py --steps 40 --no-cpu-baseline --no-random-roofline --locate-queries 0 --sharded 0 > $O/flags_$i.json 2> $O/flags_$i.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/flags_$i.json").read().strip().splitlines()[-1])
    print("[$fl] value %.1f M q/s  e2e %.1f M  index %.1f MB" % (d["value"]/1e6, d["e2e"]["value"]/1e6, d["config"]["index_hbm_bytes"]/1e6))
except Exception as ex:
    print("[$fl] failed", ex); print(open("$O/flags_$i.err").read()[-800:])
PY
done
