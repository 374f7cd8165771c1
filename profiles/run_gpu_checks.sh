#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (plain), then the ncu launch list and one
# --set full capture of the dominant kernel (profiling recipe: B200_PROFILING.md).
# Usage (from the repo root on the GPU box): bash profiles/run_gpu_checks.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/${TAG}_gpu.csv 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu.log
tail -5 $OUT/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $OUT/${TAG}_smoke.log
tail -3 $OUT/${TAG}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
tail -5 $OUT/${TAG}_bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_ref.json 2> $OUT/${TAG}_ref.err; echo "reference arm exit $?"
# launch list of the bench's own timed path (CUDA graph of the multi-batch call) and of the one-call-per-step path
SMALL="python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline"
timeout 300 $SMALL > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 $SMALL > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_states_warp -s 30 -c 2 \
    -f -o $OUT/${TAG}_k_states $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
