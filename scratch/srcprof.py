"""ncu source-page summary: python scratch/srcprof.py X.ncu-rep [top]  -> hottest source lines by stall samples + key raw metrics."""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = None; out = []
for r in rows:
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr is None or not r or not r[0].strip().isdigit(): continue
    iE = hdr.index('Instructions Executed'); iS = hdr.index('# Samples'); iT = hdr.index('Thread Instructions Executed')
    try: n = int(r[iE]); s = int(r[iS]); t = int(r[iT])
    except ValueError: continue
    out.append((int(r[0]), n, s, r[1], t))
tot = sum(o[1] for o in out); ts = sum(o[2] for o in out); tt = sum(o[4] for o in out)
print('warp instr', tot, 'samples', ts, 'avg active threads', round(tt / max(tot, 1), 1))
for ln, n, s, src, t in sorted(out, key=lambda o: -o[2])[:top]:
    print(f"{ln:4d} inst {100*n/tot:5.1f}% samp {100*s/ts:5.1f}% thr/inst {t/max(n,1):5.1f}  {src[:110]}")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); h = rows[0]; r = rows[-1]
for w in ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
          'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
          'l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum', 'lts__t_sectors_op_atom.sum', 'lts__t_sectors_op_red.sum',
          'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio','smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio']:
    for i, c in enumerate(h):
        if c == w: print(f'  {w:85s} {r[i]}')
