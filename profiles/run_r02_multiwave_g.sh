P="timeout 120 python profiles/one_kernel.py c4_fused reps=8"
for n in 20000 40000; do for g in 0 12 14 18 20; do $P n=$n poly_traj_per_warp=$g poly_blocks_per_sm=5; done; done > gpurun_out/r02_multiwave_g.log 2>&1
cat gpurun_out/r02_multiwave_g.log
