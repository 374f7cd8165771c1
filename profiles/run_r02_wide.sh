#!/bin/bash
# K3b with 64-register threads and 4 resident blocks per SM (knob poly_blocks_per_sm=4) against the default 48 / 5.
TAG=${1:-r02_wide}
OUT=gpurun_out
mkdir -p $OUT
EPP_TUNE=poly_blocks_per_sm=4 timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc" > $OUT/${TAG}_pytest_wide.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_wide.log
tail -3 $OUT/${TAG}_pytest_wide.log
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
{
$P
$P poly_blocks_per_sm=4
$P poly_blocks_per_sm=4 poly_traj_per_warp=16
$P poly_blocks_per_sm=4 poly_traj_per_warp=14
$P poly_blocks_per_sm=4 poly_ring=6
$P poly_blocks_per_sm=4 poly_wait=1
$P
$P poly_blocks_per_sm=4
$P n=20000
$P n=20000 poly_traj_per_warp=16
$P n=20000 poly_blocks_per_sm=4
$P n=20000 poly_blocks_per_sm=4 poly_traj_per_warp=16
$P n=40000
$P n=40000 poly_blocks_per_sm=4
$P n=2000
$P n=2000 poly_blocks_per_sm=4
} > $OUT/${TAG}_ab_poly.log 2>&1
cat $OUT/${TAG}_ab_poly.log
