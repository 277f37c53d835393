#!/bin/bash
# e2e (host-pointer C-ABI call) against the pipeline's chunk size.
O=gpurun_out; mkdir -p $O
for b in "$@"; do
  RS_B200_CHUNK=$b timeout 300 python bench.py --no-cpu-baseline --no-random-roofline --locate-queries 0 > $O/chunk_$b.json 2> $O/chunk_$b.err
  python - <<PY
import json
try:
    d=json.load(open("$O/chunk_$b.json"))
    print("CHUNK=$b value %.1f M q/s  e2e %.1f M" % (d["value"]/1e6, d["e2e"]["value"]/1e6))
except Exception as ex:
    print("CHUNK=$b failed", ex); print(open("$O/chunk_$b.err").read()[-1500:])
PY
done
