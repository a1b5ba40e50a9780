#!/usr/bin/env python
"""Turn gpurun_out/launches.csv and an .ncu-rep into the committed summaries under profiles/."""
import collections
import csv
import subprocess
import sys

METRICS = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
           'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
           'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
           'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
           'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
           'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
           'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
           'smsp__cycles_active.avg', 'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio',
           'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
           'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio']


def launches(path, out):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
    hdr, data = rows[hi], rows[hi + 1:]
    ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.defaultdict(list)
    for r in data:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(',', ''))
        v = v / 1e3 if r[ui] == 'ns' else (v * 1e3 if r[ui] == 'ms' else v)
        agg[r[ki].split('(')[0]].append(v)
    tot = sum(sum(v) for v in agg.values())
    with open(out, 'w') as f:
        f.write('| kernel | launches | total us | mean us | share |\n|---|---|---|---|---|\n')
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            f.write(f'| `{k}` | {len(v)} | {sum(v):.1f} | {sum(v)/len(v):.1f} | {100*sum(v)/tot:.1f}% |\n')
        f.write(f'\ntotal {tot:.1f} us over {sum(len(v) for v in agg.values())} launches (ncu-serialised, cold caches: compare shares)\n')


def raw(rep, out):
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, 'w') as f:
        for r in rows[2:]:
            f.write(f"### {r[idx['Kernel Name']]}\n\n| metric | value | unit |\n|---|---|---|\n")
            for m in METRICS:
                if m in idx:
                    f.write(f'| {m} | {r[idx[m]]} | {units[idx[m]]} |\n')
            f.write('\n')


if __name__ == '__main__':
    if sys.argv[1] == 'launches':
        launches(sys.argv[2], sys.argv[3])
    else:
        raw(sys.argv[2], sys.argv[3])
