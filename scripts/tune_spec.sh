#!/bin/bash
# scheduler knobs after the speculative-evaluation changes (fewer evaluations, longer control tasks), one box: ms per step of the default bench
for cfg in "" "SDSM_CTRL_CTAS=56" "SDSM_CTRL_CTAS=64" "SDSM_CTRL_CTAS=80" "SDSM_CTRL_CTAS=100" "SDSM_CTRL_SHARE=1.3" "SDSM_CTRL_CTAS=64 SDSM_HI_EVALS=150" "SDSM_HI_FRAC=0.06" "SDSM_SOLO_CELLS=16" "SDSM_TASK_SMALL_CELLS=1024" ""; do
  ms=$(env $cfg timeout 300 python bench.py --steps 4 --warmup 4 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(d['ms_per_step'],2), d['stats']['rank0_step_ms'], d['solver']['control_cta_time_share'], d['solver']['evaluation_warp_time_share'])")
  echo "[$cfg] $ms"
done
