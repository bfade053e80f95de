"""Summarise an .ncu-rep (read on the CPU box): key metrics, opcode mix, hottest source lines."""
import collections, csv, io, re, subprocess, sys
rep = sys.argv[1]
def run(*a):
    return subprocess.run(["ncu", "-i", rep, *a], capture_output=True, text=True).stdout
raw = list(csv.reader(io.StringIO(run("--page", "raw", "--csv"))))
hdr, units = raw[0], raw[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__grid_size',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts.sum',
        'smsp__warps_eligible.avg.per_cycle_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__cycles_elapsed.max', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_uniform.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct']
for r in raw[2:]:
    print('---', r[hdr.index('Kernel Name')][:60])
    for k in keys:
        if k in hdr:
            print(f'  {k:82s} {r[hdr.index(k)]:>16s} {units[hdr.index(k)]}')
cs = list(csv.reader(io.StringIO(run("--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-count", "1"))))
lines, ops, tot = [], collections.Counter(), 0
seen = set()
for r in cs[3:]:
    if not r: continue
    if r[0].strip().isdigit():
        try: lines.append((int(r[7]), int(r[6]) if r[6].isdigit() else 0, int(r[0]), r[1].strip()[:100]))
        except Exception: pass
    elif len(r) > 7 and r[2].startswith('0x') and r[2] not in seen:
        seen.add(r[2])
        m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[3])
        try: n = int(r[7])
        except Exception: continue
        ops[m.group(2).split('.')[0] if m else '?'] += n; tot += n
print('total warp instructions', tot)
print('opcode mix:', ', '.join(f'{o} {100*n/tot:.1f}%' for o, n in ops.most_common(16)))
lt = sum(l[0] for l in lines)
for n, smp, ln, src in sorted(lines, key=lambda x: -x[0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print(f'{n:10d} {100*n/lt:5.1f}% samp={smp:5d}  L{ln:4d} {src}')
