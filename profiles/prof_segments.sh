OUT=gpurun_out; mkdir -p $OUT
for v in 0 1; do
SMALL="python bench.py --workload segments --steps 32 --warmup 3 --no-graph --no-extras --no-cpu-baseline --tune segments_variant=$v"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_segments -s 5 -c 2 -f -o $OUT/r01_w_k_segments_v$v $SMALL > $OUT/r01_w_ncu_seg_v$v.log 2>&1
echo "ncu v$v exit $?"
done
