OUT=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/r02_last_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/r02_last_pytest_gpu.log
tail -3 $OUT/r02_last_pytest_gpu.log
P="timeout 120 python profiles/one_kernel.py c4_fused reps=8"
for n in 500 2000 5000 10000 12000 20000 40000; do $P n=$n; done > $OUT/r02_shapes_defaults.log 2>&1
cat $OUT/r02_shapes_defaults.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/r02_last_bench.json 2> $OUT/r02_last_bench.err; echo "bench exit $?"
