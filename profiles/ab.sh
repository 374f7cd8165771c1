#!/bin/bash
# quick A/B timings of the round-2 knobs (one_kernel.py); usage: bash profiles/ab.sh > gpurun_out/ab.log
P="python profiles/one_kernel.py"
T2="states_warp_blocks_per_sm=2 states_warp_threads=512"
$P c2_states n=16777216; $P c2_states n=16777216 $T2
$P w256_states; $P w256_states $T2; $P w256_states n=16777216; $P w256_states n=16777216 $T2
$P c4_pre; $P c4_pre $T2
$P c5_states; $P c5_states $T2
$P c4_fused; $P c4_fused poly_threads=512 poly_traj_per_warp=32; $P c4_fused poly_threads=512 poly_traj_per_warp=16; $P c4_fused poly_traj_per_warp=8
$P c3_segments
