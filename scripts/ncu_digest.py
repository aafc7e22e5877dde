#!/usr/bin/env python3
"""Digest of an ncu raw/source CSV pair: key metrics, opcode mix, basic-block profile."""
import csv, re, sys
from collections import Counter
raw, src = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ['l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum', 'lts__t_sectors_srcunit_tex_op_read.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__grid_size', 'launch__block_size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__sass_average_branch_targets_threads_uniform.pct',
        'smsp__inst_executed.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__warps_eligible.avg.per_cycle_active',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio']
for h, u, v in zip(hdr, units, vals):
    if h in keep:
        print(f'{h} [{u}] = {v}')
rows = list(csv.reader(open(src)))
hdr = rows[1]
ia, ie, it = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Avg. Threads Executed')
ops, tot, data = Counter(), 0, []
for r in rows[2:]:
    try:
        n = int(r[ie])
    except Exception:
        continue
    sass = r[ia].strip()
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)', sass)
    op = m.group(2).split('.')[0] if m else sass[:10]
    ops[op] += n
    tot += n
    data.append((n, r[it], sass))
print('total warp instructions', tot)
for op, n in ops.most_common(24):
    print(f'  {op:10s} {n/1e9:9.3f}G {100*n/tot:5.1f}%')
prev, start, cnt = None, 0, 0
for k, (n, t, s) in enumerate(data + [(None, None, None)]):
    key = (n, t)
    if key != prev:
        if prev is not None and prev[0] and prev[0] > tot / 5000:
            print(f'  block {start:4d}-{k-1:4d} len={k-start:3d} exec={prev[0]/1e9:8.4f}G thr={prev[1]}')
        prev, start = key, k
if len(sys.argv) > 3:
    with open(sys.argv[3], 'w') as f:
        for k, (n, t, s) in enumerate(data):
            f.write(f'{k:4d} {n:>12d} {t:>5s} {s}\n')
