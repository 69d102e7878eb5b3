#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench, then the ncu launch list and a full capture of the step kernel.
# Usage (from the repo root, under gpurun): bash scripts/gpu_check.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1
nproc >> $OUT/${TAG}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?" | tee -a $OUT/${TAG}_smoke.log
timeout 900 python bench.py --steps 2000 --warmup 20 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
tail -c 3000 $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err
timeout 900 python bench.py --rows > $OUT/${TAG}_kernels.jsonl 2> $OUT/${TAG}_kernels.err; echo "kernels exit $?"
BCMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $BCMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv $BCMD > $OUT/${TAG}_ncu1.log 2>&1
timeout 300 $BCMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sokoban_tile -s 4 -c 2 -f -o $OUT/${TAG}_step $BCMD > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu exit $?"
ls -la $OUT
