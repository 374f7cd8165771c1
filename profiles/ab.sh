#!/bin/bash
# quick A/B timings of the round-2 knobs (one_kernel.py); usage: bash profiles/ab.sh > gpurun_out/ab.log
P="python profiles/one_kernel.py"
$P c2_states; $P c2_states n=16777216
$P w256_states; $P w256_states bits_cmask=0
$P c3_segments; $P c3_segments bits_range_d=0; $P c3_segments bits_range_d=8; $P c3_segments segments_variant=3; $P c3_segments segments_warp_slots=32
$P c4_fused; $P c4_fused poly_traj_per_warp=32; $P c4_fused poly_traj_per_warp=8; $P c4_pre
$P c5_states; $P c5_states grid_occ_kb=40; $P c5_states grid_occ_kb=56; $P c5_states grid_occ_kb=20; $P c5_states states_warp_log2_slots=5
$P c5_edges; $P c5_edges grid_occ_kb=40;  $P c5_edges grid_occ_kb=20; $P c5_edges segments_warp_slots=32
