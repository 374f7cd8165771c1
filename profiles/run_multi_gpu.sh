#!/bin/bash
# One `gpurun --gpus N` call: the default bench (config 2, weak scaling) and config 5 (strong scaling)
# at 1 and N ranks, plus the reference arm launched the way the driver launches it.
# Usage (repo root on the GPU box): bash profiles/run_multi_gpu.sh N [tag]
N=${1:-2}
TAG=${2:-r1}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1
timeout 600 python bench.py --steps 2000 --warmup 16 --no-extras > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err; echo "bench n1 exit $?"
cat $OUT/${TAG}_bench_n1.json
timeout 600 $TR bench.py --gpus $N --steps 2000 --warmup 16 > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err; echo "bench n$N exit $?"
cat $OUT/${TAG}_bench_n$N.json; tail -3 $OUT/${TAG}_bench_n$N.err
timeout 900 python bench.py --workload config5 > $OUT/${TAG}_c5_n1.json 2> $OUT/${TAG}_c5_n1.err; echo "config5 n1 exit $?"
cat $OUT/${TAG}_c5_n1.json; tail -3 $OUT/${TAG}_c5_n1.err
timeout 900 $TR bench.py --gpus $N --workload config5 --no-cpu-baseline > $OUT/${TAG}_c5_n$N.json 2> $OUT/${TAG}_c5_n$N.err; echo "config5 n$N exit $?"
cat $OUT/${TAG}_c5_n$N.json; tail -3 $OUT/${TAG}_c5_n$N.err
timeout 600 $TR bench.py --impl reference --gpus $N --steps 20 --warmup 3 > $OUT/${TAG}_ref_n$N.json 2> $OUT/${TAG}_ref_n$N.err; echo "reference arm exit $?"
cat $OUT/${TAG}_ref_n$N.json
