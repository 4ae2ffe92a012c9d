#!/bin/bash
for cfg in "" "SDSM_CTRL_CTAS=250" "SDSM_CTRL_CTAS=280" "SDSM_CTRL_CTAS=290"; do
  ms=$(env $cfg timeout 600 python bench.py --workload tiled4k --field 2048 --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(d['ms_per_step'],1), round(d['value']), d['solver']['control_cta_time_share'], d['solver']['evaluation_warp_time_share'])")
  echo "[$cfg] $ms"
done
