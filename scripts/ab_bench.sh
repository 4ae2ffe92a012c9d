#!/bin/bash
# A/B of library builds on one box: scripts/ab_bench.sh ab_tmp/libA.so ab_tmp/libB.so ...   (each: bench ms/step twice)
for rep in 1 2; do
  for lib in "$@"; do
    ms=$(SDSM_LIB=$PWD/$lib timeout 300 python bench.py --steps 4 --warmup 4 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(d['ms_per_step'],2), d['stats']['rank0_step_ms'])")
    echo "$lib rep$rep ms_per_step $ms"
  done
done
