P="timeout 120 python profiles/one_kernel.py c4_fused reps=8"
for n in 10000 20000 40000 2000; do for t in 256 224 192; do $P n=$n poly_threads=$t; done; done > gpurun_out/r02_slim_ab.log 2>&1
$P n=10000 poly_threads=192 poly_ring=6 >> gpurun_out/r02_slim_ab.log 2>&1
$P n=10000 poly_threads=224 poly_ring=6 >> gpurun_out/r02_slim_ab.log 2>&1
cat gpurun_out/r02_slim_ab.log
EPP_TUNE=poly_threads=192 timeout 600 python -m pytest tests -m gpu -x -q -k "poly or no_alloc" 2>&1 | tail -1
EPP_TUNE=poly_threads=224 timeout 600 python -m pytest tests -m gpu -x -q -k "poly or no_alloc" 2>&1 | tail -1
