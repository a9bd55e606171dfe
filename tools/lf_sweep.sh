#!/bin/bash
# Sweep of the long-filter super-block plan: blocks per launch triple x lanes (bench.py --workload longfir).
for lanes in 2 3 4; do for cb in 2 3 4; do
  r=$(PEBBLE_LF_LANES=$lanes PEBBLE_LF_SUPER_CHUNK_BLOCKS=$cb timeout 300 python bench.py --workload longfir --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['frac'], d['roofline'].get('traffic'))")
  echo "lanes=$lanes super_chunk=$cb $r"
done; done
