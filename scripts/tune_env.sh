#!/bin/bash
# scripts/tune_env.sh "VAR=val VAR2=val" ...: bench's device/e2e numbers under each environment setting.
O=gpurun_out; mkdir -p $O; i=0
for ev in "$@"; do
  i=$((i+1))
  env $ev timeout 300 python bench.py --steps 60 --no-cpu-baseline --no-random-roofline --locate-queries 0 --sharded 0 > $O/env_$i.json 2> $O/env_$i.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/env_$i.json").read().strip().splitlines()[-1])
    print("[$ev] value %.1f M q/s  e2e %.1f M" % (d["value"]/1e6, d["e2e"]["value"]/1e6))
except Exception as ex:
    print("[$ev] failed", ex); print(open("$O/env_$i.err").read()[-800:])
PY
done
