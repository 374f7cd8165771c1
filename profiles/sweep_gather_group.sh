#!/bin/bash
# N-rank sweep of the fused exchange's send granularity (tuning knob gather_group_chunks).
N=${1:-8}; TAG=${2:-r1}; shift; shift
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"
for g in "$@"; do
  echo "== gather_group_chunks=$g"
  timeout 300 $TR bench.py --gpus $N --steps 2000 --warmup 16 --no-extras --no-cpu-baseline --tune gather_group_chunks=$g 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'], l['config']['exchange_verified'])"
done 2>&1 | tee $OUT/${TAG}_sweep_gather_group_n$N.log
