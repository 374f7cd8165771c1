for t in "" "states_per_thread=2" "states_per_thread=2 states_warp_threads=768" "states_per_thread=8"; do
  args=""; for kv in $t; do args="$args --tune $kv"; done
  echo "== $t"; timeout 300 python bench.py --steps 2000 --warmup 16 --no-extras --no-cpu-baseline $args 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step']*1e3, 'us', d['roofline']['frac'], d['config']['mask_popcount_slot0'])"
done
