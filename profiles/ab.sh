#!/bin/bash
# quick A/B timings of the round-2 knobs (one_kernel.py); usage: bash profiles/ab.sh > gpurun_out/ab.log
P="python profiles/one_kernel.py"
$P c2_states; $P c2_states states_table_copies=8; $P c2_states states_table_copies=4; $P c2_states n=16777216; $P c2_states n=16777216 states_table_copies=8
$P w256_states; $P w256_states bits_cmask=0; $P w256_states morton_sort=0 bits_cmask=0
$P c3_segments; $P c3_segments bits_cmask=0; $P c3_segments morton_sort=0 bits_cmask=0; $P c3_segments segments_variant=3; $P c3_segments segments_warp_threads=768; $P c3_segments segments_warp_threads=512
$P c4_fused; $P c4_fused poly_traj_per_warp=32; $P c4_fused poly_traj_per_warp=8; $P c4_fused bits_cmask=0; $P c4_pre
$P c5_states; $P c5_states grid_cell_recs=0; $P c5_states morton_sort=0; $P c5_states grid_occ_kb=40; $P c5_states grid_occ_kb=110
$P c5_edges; $P c5_edges grid_cell_recs=0; $P c5_edges segments_warp_threads=768; $P c5_edges segments_variant=3
python - <<'PY'
import epp_b200, torch
from epp_b200 import worlds
config = epp_b200.load_config()
for kind in ("track", "w256", "w10k"):
    g, o, lo, hi = {"track": lambda: worlds.track_world(config), "w256": worlds.world_256, "w10k": worlds.world_10k}[kind]()
    w = epp_b200.World(config); w.parse_gates_and_obstacles(g, o)
    print(kind, "single isValid us/call (mailbox):", w.single_latency_us(5000))
    epp_b200.set_tuning("single_mailbox", 0); print(kind, "  launch+sync:", w.single_latency_us(300)); epp_b200.set_tuning("single_mailbox", 1)
PY
