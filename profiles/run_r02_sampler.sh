#!/bin/bash
# New sampler emission loop of K3b: parity (all sweep tests), then same-box A/B against the previous library.
OUT=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc or planner" > $OUT/r02_sampler_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/r02_sampler_pytest.log
tail -3 $OUT/r02_sampler_pytest.log
EPP_TUNE=poly_blocks_per_sm=5 timeout 900 python -m pytest tests -m gpu -x -q -k "poly or no_alloc" 2>&1 | tail -1
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
PREV=profiles/_prev/libepp_b200_prev.so
{
for n in 10000 2000 5000 20000 40000 500; do
  echo "new:"; $P n=$n
  echo "prev:"; EPP_B200_LIB=$PREV $P n=$n
done
echo "new:"; $P; echo "prev:"; EPP_B200_LIB=$PREV $P
} > $OUT/r02_sampler_ab.log 2>&1
cat $OUT/r02_sampler_ab.log
