#!/bin/bash
# One gpurun call at the end of round 2: the allocation-free entry point's test alone in a fresh process, the whole GPU
# suite, the polynomial-sweep tests with the opt-in K3b knobs, smoke, and A/B timings of those knobs on config 4.
TAG=${1:-r02_fin}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/${TAG}_gpu.csv 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/${TAG}_pytest_no_alloc_alone.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_no_alloc_alone.log
tail -4 $OUT/${TAG}_pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu.log
tail -4 $OUT/${TAG}_pytest_gpu.log
EPP_TUNE=poly_wait=1,poly_balance=1 timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc" > $OUT/${TAG}_pytest_poly_knobs.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_poly_knobs.log
tail -4 $OUT/${TAG}_pytest_poly_knobs.log
timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $OUT/${TAG}_smoke.log
tail -2 $OUT/${TAG}_smoke.log
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
{
$P
$P poly_wait=1
$P poly_balance=1
$P poly_wait=1 poly_balance=1
$P poly_wait=1 poly_balance=1 poly_ring=6
$P poly_wait=1 poly_balance=1 poly_ring=8
$P poly_wait=1 poly_traj_per_warp=12
$P poly_wait=1 poly_traj_per_warp=14
$P poly_wait=1 poly_traj_per_warp=18
$P poly_wait=1 poly_traj_per_warp=28
$P poly_wait=1 poly_balance=1 n=2000
$P n=2000
$P poly_wait=1 poly_balance=1 n=40000
$P n=40000
} > $OUT/${TAG}_ab_poly.log 2>&1
cat $OUT/${TAG}_ab_poly.log
