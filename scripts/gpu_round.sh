#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, the ncu launch list and one --set full capture of K1.
# Usage (from the repo root): gpurun --timeout 1700 -- 'bash scripts/gpu_round.sh TAG'
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $O/gpu_$TAG.txt 2>&1
nproc >> $O/gpu_$TAG.txt
if [ "${SKIP_TESTS:-0}" != 1 ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$TAG.log
  tail -3 $O/pytest_gpu_$TAG.log
fi
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke_$TAG.log
timeout 400 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench rc=$?"; cat $O/bench_$TAG.json
if [ "${SKIP_REF:-0}" != 1 ]; then
  timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "ref rc=$?"; cat $O/bench_ref_$TAG.json
fi
# launch list of the same command (serialised, cold-cache times: shares must agree, not absolutes)
timeout 400 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file $O/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-random-roofline --sharded 0 \
  > $O/ncu_launches_$TAG.log 2>&1; echo "ncu list rc=$?"
# one full capture of the dominant kernel
timeout 500 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:k_find_interval -s 3 -c 1 \
  -o $O/k1_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-random-roofline --locate-queries 0 --sharded 0 \
  > $O/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ncu -i $O/k1_$TAG.ncu-rep --page raw --csv > $O/k1_$TAG.raw.csv 2>/dev/null
python profiles/ncu_summary.py $O/k1_$TAG.raw.csv > $O/k1_$TAG.summary.txt 2>&1; head -40 $O/k1_$TAG.summary.txt
