#!/bin/bash
# A/B of library builds on one box: scripts/ab_bench.sh ab_tmp/libA.so ab_tmp/libB.so ...   (each: bench ms/step twice, sweep once)
for rep in 1 2; do
  for lib in "$@"; do
    ms=$(SDSM_LIB=$PWD/$lib timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2), d['solver']['evaluation_warp_time_share'])")
    echo "$lib rep$rep ms_per_step $ms"
  done
done
for lib in "$@"; do
  echo "$lib sweep:"; SDSM_LIB=$PWD/$lib timeout 300 python scripts/sweep_bench.py 2>&1 | head -3
done
