#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): per-kernel key metrics + top stall PCs."""
import csv, subprocess, sys, io
rep = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof.ncu-rep"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct',
        'sm__cycles_elapsed.max', 'launch__grid_size', 'launch__block_size']
ki = hdr.index('Kernel Name')
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    print("=====", r[ki][:90])
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"  {w:75s} {r[i]:>16s} {units[i]}")
    st = []
    for i, h in enumerate(hdr):
        if h.startswith('smsp__pcsamp_warps_issue_stalled_') and 'not_issued' not in h:
            try:
                st.append((float(r[i].replace(',', '')), h.replace('smsp__pcsamp_warps_issue_stalled_', '')))
            except ValueError:
                pass
    tot = sum(v for v, _ in st) or 1
    print("  stalls: " + ", ".join(f"{n} {100*v/tot:.0f}%" for v, n in sorted(st, reverse=True)[:9]))
