"""Step time of the GOWT1-shaped workload per image seed (device-resident, one launch per image)."""
import sys
sys.path.insert(0, '.')
import numpy as np, torch
from workloads import synth
from superdsm_b200.objects import get_engine
eng = get_engine(0)
for seed in range(1, 9):
    w = synth.make_workload('gowt1', oversegment=3, max_set_size=4, seed=seed)
    eng.set_image(w['y'], w['y_mask'], w['atoms'], w['dsm_cfg']['background_margin'])
    eng.set_params(w['dsm_cfg'])
    h = eng.prepare(w['footprints'])
    stream = torch.cuda.current_stream()
    eng.run_prepared(h, stream); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); eng.run_prepared(h, stream); eng.run_prepared(h, stream); e1.record(stream); torch.cuda.synchronize()
    n = h['cands']['n']
    info = h['out']['info'][:32 * h['n']].cpu().numpy().view(np.int32).reshape(-1, 8)
    ft = eng.finish_times_us(h['ws'], h['n'], h['plan'])
    k = int(np.argmax(ft))
    print(f'seed {seed}: {h["n"]} candidates, {int(n.sum())} px, {e0.elapsed_time(e1) / 2:.1f} ms/step; last finisher cand {k} n {n[k]} m {info[k,0]} it {info[k,1:4].tolist()} nf {info[k,6]} nh {info[k,7]}; max n {n.max()}')
    del h
    torch.cuda.empty_cache()
