#!/bin/bash
OUT=gpurun_out
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
{ $P; $P poly_prefetch=1; $P; $P poly_prefetch=1; $P n=40000; $P n=40000 poly_prefetch=1; $P n=20000; $P n=20000 poly_prefetch=1; } > $OUT/r02_prefetch_ab.log 2>&1
cat $OUT/r02_prefetch_ab.log
EPP_TUNE=poly_prefetch=1 timeout 600 python -m pytest tests -m gpu -x -q -k "poly or no_alloc" 2>&1 | tail -2
