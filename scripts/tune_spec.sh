#!/bin/bash
# A/B of speculative backtracking and the scheduler knobs it shifts, on one box: ms per step of the default bench (4-image pool)
for cfg in "SDSM_SPEC_MAX=1" "" "SDSM_SPEC_SOLO=0" "SDSM_SPEC_PRIO=1" "SDSM_CTRL_CTAS=36" "SDSM_CTRL_CTAS=28" "SDSM_CTRL_SHARE=0.7" "SDSM_CTRL_CTAS=36 SDSM_CTRL_SHARE=0.7" "SDSM_SOLO_CELLS=64" "SDSM_SPEC_RETRY=2" "SDSM_SPEC_MAX=1"; do
  ms=$(env $cfg timeout 300 python bench.py --steps 4 --warmup 4 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(d['ms_per_step'],2), d['stats']['rank0_step_ms'], d['solver']['control_cta_time_share'], d['solver']['evaluation_warp_time_share'])")
  echo "[$cfg] $ms"
done
