#!/usr/bin/env python3
"""Summarise an .ncu-rep: per-launch metrics (raw page) and, per kernel, the hottest SASS lines (source page)."""
import csv, subprocess, sys, io

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'smsp__cycles_active.avg', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sectors.sum',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']

def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    ik = hdr.index('Kernel Name')
    idx = [(w, hdr.index(w)) for w in WANT if w in hdr]
    for r in rows[2:]:
        print(r[ik][:28].ljust(28), ' '.join(f"{w.split('.')[0].split('__')[-1][:18]}={r[i][:10]}" for w, i in idx))

def source(rep, kernel, top=25):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', kernel], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = None; data = []; launches = 0
    for r in rows:
        if r and r[0] == 'Address':
            hdr = r; launches += 1
            if launches > 1: break
            continue
        if hdr is None or len(r) < len(hdr) - 2: continue
        try:
            data.append((int(r[hdr.index('# Samples')] or 0), int(r[hdr.index('Instructions Executed')] or 0), r[hdr.index('Source')].strip(), len(data)))
        except ValueError:
            pass
    tot = sum(d[0] for d in data)
    print(f"== {kernel}: {tot} samples, {sum(d[1] for d in data)} warp instructions, {len(data)} SASS lines")
    for d in sorted(data, key=lambda x: -x[0])[:top]:
        print(f"  line {d[3]:5d} samples {d[0]:6d} ({100.0 * d[0] / max(tot, 1):4.1f}%) execs {d[1]:9d}  {d[2][:90]}")

if __name__ == '__main__':
    rep = sys.argv[1]
    raw(rep)
    for k in sys.argv[2:]:
        source(rep, k)
