#!/bin/bash
# One `gpurun --gpus G` call: the default bench (config 2, weak scaling, fused mask exchange) at N = 1, 2, 4, .. G,
# config 5 (strong scaling) at 1 and G ranks, the NCCL form at G for comparison.
# Usage (repo root on the GPU box): bash profiles/run_scaling.sh G [tag]
G=${1:-8}
TAG=${2:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1
run() { # N, extra args..., output name
  local n=$1; shift; local name=$1; shift
  if [ "$n" = 1 ]; then timeout 900 python bench.py "$@" > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 \
         bench.py --gpus $n "$@" > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err; fi
  echo "$name exit $?"; python - <<PY
import json
try:
    l = json.loads([x for x in open("$OUT/${TAG}_$name.json") if x.startswith("{")][-1])
    print("  n_gpus", l["n_gpus"], "value %.4g" % l["value"], "ms/step %.5f" % l["ms_per_step"], "e2e %.4g" % l["e2e"]["value"],
          l["config"].get("exchange", ""), l["config"].get("exchange_verified", ""))
except Exception as ex:
    print("  no line:", ex)
PY
}
N=1
while [ $N -le $G ]; do
  run $N bench_n$N --steps 2000 --warmup 16 --no-extras --no-cpu-baseline
  N=$((N * 2))
done
run $G bench_nccl_n$G --steps 2000 --warmup 16 --no-extras --no-cpu-baseline --exchange nccl
run 1 c5_n1 --workload config5 --no-cpu-baseline
run $G c5_n$G --workload config5 --no-cpu-baseline
run $G c5_nccl_n$G --workload config5 --no-cpu-baseline --exchange nccl
run $G ref_n$G --impl reference --steps 20 --warmup 3
