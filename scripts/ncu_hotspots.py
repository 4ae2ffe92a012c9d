"""Summaries of an `ncu --set full --import-source on` report for profiles/: python scripts/ncu_hotspots.py <rep> <tag> [title]
writes profiles/<tag>_ncu_summary.csv (selected raw metrics) and profiles/<tag>_hotspots.txt (stall samples by source line)."""
import csv, subprocess, sys, io, os
rep, tag = sys.argv[1], sys.argv[2]
title = sys.argv[3] if len(sys.argv) > 3 else tag
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes_read.sum.per_second', 'dram__bytes_write.sum.per_second',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__inst_executed.avg.per_cycle_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__pcsamp_sample_count', 'sm__icc_request_hit_rate.pct',
        'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_fp64.sum',
        'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum']
want += [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
with open(f'profiles/{tag}_ncu_summary.csv', 'w') as f:
    f.write(f'# {title}\nmetric,unit,value\n')
    for h, u, v in zip(hdr, units, vals):
        if h in want: f.write(f'{h},{u},{v}\n')
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
cur = None; agg = {}; tot = 0; totl = 0; il = None
for row in csv.reader(io.StringIO(src)):
    if len(row) >= 2 and row[0] == 'File Path': cur = row[1].split('/')[-1]; continue
    if len(row) > 3 and row[0] == 'Line No': il = row.index('stall_long_sb'); continue
    if len(row) < 8 or row[0] == '': continue
    try: ln = int(row[0]); s = int(row[4]); inst = int(row[7]); l = int(row[il])
    except Exception: continue
    agg[(cur, ln)] = (s, inst, l, row[1][:95]); tot += s; totl += l
with open(f'profiles/{tag}_hotspots.txt', 'w') as f:
    f.write(f'# {title}: warp-stall samples by source line (inlined call sites are counted at both levels); long_sb = samples stalled on a global/local load\n')
    f.write(f'total samples {tot}, of which long-scoreboard {totl}\n')
    for (fl, l), (s, i, ls, t) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:60]:
        f.write(f'{fl}:{l:4d} {100 * s / tot:5.1f}% long_sb {100 * ls / tot:5.1f}% warp-inst {i:>11d}  {t}\n')
print(open(f'profiles/{tag}_hotspots.txt').read()[:6000])
