#!/bin/bash
for wl in bbbc039 nih3t3 tiled; do
 for cfg in "SDSM_CTRL_SHARE=0.65" "SDSM_CTRL_SHARE=0.85" "SDSM_CTRL_SHARE=0.95"; do
  ms=$(env $cfg timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(d['ms_per_step'],1), round(d['value']), d['solver']['control_cta_time_share'], d['solver']['evaluation_warp_time_share'])")
  echo "$wl [$cfg] $ms"
 done
done
