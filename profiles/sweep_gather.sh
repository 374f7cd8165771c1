#!/bin/bash
# N-rank sweep of the fused exchange's parts (profiling only): gather_debug 1 no wait/unpack, 2 no peer stores.
N=${1:-2}
TAG=${2:-r1}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
B="bench.py --gpus $N --steps 2000 --warmup 16 --no-extras --no-cpu-baseline"
for d in 0 3 7; do
  echo "== gather_debug=$d"
  timeout 300 $TR $B --tune gather_debug=$d --no-verify 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'], l['config']['exchange_verified'])"
done 2>&1 | tee $OUT/${TAG}_sweep_gather_n$N.log
echo "== states_pdl=0"
timeout 300 $TR $B --tune states_pdl=0 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'], l['config']['exchange_verified'])" | tee -a $OUT/${TAG}_sweep_gather_n$N.log
echo "== no exchange at all (plain kernel per rank)"
timeout 300 $TR $B --exchange none --no-verify 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'], l['config']['launch'])" | tee -a $OUT/${TAG}_sweep_gather_n$N.log
echo "== no exchange, plain kernel writing into IPC-exported memory"
timeout 300 $TR $B --exchange none-ipc --no-verify 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'], l['config']['launch'])" | tee -a $OUT/${TAG}_sweep_gather_n$N.log
echo "== nccl"
timeout 300 $TR $B --exchange nccl 2>/dev/null | python -c "import sys,json; l=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print(l['ms_per_step']*1e3,'us/step', l['value'])" | tee -a $OUT/${TAG}_sweep_gather_n$N.log
