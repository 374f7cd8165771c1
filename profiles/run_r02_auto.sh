timeout 900 python -m pytest tests -m gpu -x -q -k "poly or traj or no_alloc" 2>&1 | tail -2
P="timeout 120 python profiles/one_kernel.py c4_fused reps=10"
{ $P; $P n=7500; $P n=7500 poly_blocks_per_sm=5 poly_balance=0; $P n=12000; $P n=12000 poly_traj_per_warp=16; $P n=20000; $P n=5000; } > gpurun_out/r02_auto_ab_poly.log 2>&1
cat gpurun_out/r02_auto_ab_poly.log
bash profiles/prof_kernels.sh r02_q c3_segments:k_segments_warp
