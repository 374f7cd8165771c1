#!/bin/bash
# Segment cache of the polynomial sweep (knob poly_cache): parity tests with the knob on, then A/B timings on config 4.
TAG=${1:-r02_cache}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc" > $OUT/${TAG}_pytest_default.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_default.log
tail -4 $OUT/${TAG}_pytest_default.log
EPP_TUNE=poly_cache=1 timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc" > $OUT/${TAG}_pytest_cache.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest_cache.log
tail -4 $OUT/${TAG}_pytest_cache.log
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
{
$P
$P poly_cache=1
$P poly_cache=1 poly_ring=3
$P poly_cache=1 poly_ring=2
$P poly_cache=1 poly_traj_per_warp=12
$P poly_cache=1 poly_traj_per_warp=16
$P poly_cache=1 poly_traj_per_warp=18 poly_ring=3
$P poly_cache=1 poly_wait=1
$P n=2000
$P poly_cache=1 n=2000
$P n=40000
$P poly_cache=1 n=40000
$P poly_cache=1 n=40000 poly_ring=3
} > $OUT/${TAG}_ab_poly.log 2>&1
cat $OUT/${TAG}_ab_poly.log
