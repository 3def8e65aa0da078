#!/usr/bin/env python
"""Key metrics and the top stall lines of every kernel in an .ncu-rep (ncu -i ... --page raw/source --csv)."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','launch__grid_size','l1tex__throughput.avg.pct_of_peak_sustained_active','lts__throughput.avg.pct_of_peak_sustained_elapsed','dram__throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
idx = {h: i for i, h in enumerate(hdr)}
for d in data:
    print('====', d[idx['Kernel Name']][:60], 'id', d[idx['ID']])
    for w in want:
        if w in idx: print('  %-70s %s %s' % (w, d[idx[w]][:30], units[idx[w]]))
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", d[idx['ID']], "--launch-count", "1"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    sh = None
    for i, r in enumerate(srows):
        if r and r[0] == 'Address': sh = i; break
    if sh is None: continue
    ix = {h: i for i, h in enumerate(srows[sh])}
    def num(x):
        try: return int(x)
        except: return 0
    body = [r for r in srows[sh+1:] if len(r) > ix['# Samples'] and r[0] != 'Address']
    seen = set(); uniq = []
    for r in body:
        if r[0] in seen: continue
        seen.add(r[0]); uniq.append(r)
    tot = sum(num(r[ix['# Samples']]) for r in uniq)
    print('  samples', tot, 'warp instr', sum(num(r[ix['Instructions Executed']]) for r in uniq))
    agg = {}
    for k in ['stall_long_sb','stall_short_sb','stall_barrier','stall_math','stall_wait','stall_lg','stall_mio','stall_not_selected','stall_selected','stall_branch_resolving','stall_no_inst','stall_membar','stall_dispatch','stall_drain']:
        if k in ix: agg[k] = sum(num(r[ix[k]]) for r in uniq)
    print('  ', {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    for r in sorted(uniq, key=lambda r: -num(r[ix['# Samples']]))[:14]:
        print('   %5s %8s %-60s long %s short %s bar %s lg %s' % (r[ix['# Samples']], r[ix['Instructions Executed']], r[ix['Source']][:60], r[ix['stall_long_sb']], r[ix['stall_short_sb']], r[ix['stall_barrier']], r[ix['stall_lg']]))
