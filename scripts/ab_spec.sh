# A/B of the speculative-evaluation policy on one box: whole-workload solve times (and a bitwise comparison of all results
# with the library without speculation), then the default bench.  The baseline library is built from the commit before the
# change: git archive 5536676 superdsm_b200 include | tar -x -C /tmp/base && (cd /tmp/base && nvcc -gencode arch=compute_100a,code=sm_100a
# -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC -shared -o $REPO/superdsm_b200/libsuperdsm_b200_base.so superdsm_b200/csrc/sdsm.cu)
B=$PWD/superdsm_b200/libsuperdsm_b200_base.so
SDSM_LIB=$B python scripts/dump_results.py gpurun_out/ab_base.npz gowt1 tiled bbbc039 nih3t3 2>&1 | grep -v Warn
python scripts/dump_results.py gpurun_out/ab_new.npz gowt1 tiled bbbc039 nih3t3 2>&1 | grep -v Warn
python scripts/dump_results.py --compare gpurun_out/ab_base.npz gpurun_out/ab_new.npz | grep -v identical
for cfg in "SDSM_SPEC_MAX=1 SDSM_SPEC_HESS=0 SDSM_SPEC_SKIP=0" "" "SDSM_CTRL_CTAS=64"; do
  env $cfg timeout 300 python bench.py --steps 4 --warmup 4 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('[$cfg]', round(d['ms_per_step'],2), d['stats']['rank0_step_ms'], d['value'], d['e2e']['value'], d['solver'], d['roofline']['frac'], d['roofline']['fp64']['frac'])"
done
python -m pytest tests/test_gpu_round2.py -q -m gpu -k speculative 2>&1 | tail -3
