P="timeout 120 python profiles/one_kernel.py c5_edges reps=8"
for t in 1024 768 640 512 1024 768; do $P segments_warp_threads=$t; done > gpurun_out/r02_seg768_ab.log 2>&1
for t in 1024 768; do $P n=33554432 segments_warp_threads=$t; done >> gpurun_out/r02_seg768_ab.log 2>&1
cat gpurun_out/r02_seg768_ab.log
EPP_TUNE=segments_warp_threads=768 timeout 600 python -m pytest tests -m gpu -x -q -k "segment or edges or motion or config5 or w10k or grid" 2>&1 | tail -1
