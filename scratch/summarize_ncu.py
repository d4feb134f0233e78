"""Text summary of an `ncu -i X.ncu-rep --page raw --csv` export (+ optional --page source --csv): the counters
DESIGN.md cites, the warp-state sampling breakdown and the hottest instructions."""
import csv
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_static",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tex.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__t_requests_pipe_tex_mem_texture.sum",
        "l1tex__t_sectors_pipe_tex_mem_texture.sum", "l1tex__t_sectors_pipe_tex_mem_texture_lookup_miss.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum",
        "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio"]


def main(raw, src=None, max_kernels=4):
    rows = list(csv.reader(open(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:2 + max_kernels]:
        d = dict(zip(hdr, zip(units, r)))
        print("kernel:", d.get("Kernel Name", ("", "?"))[1][:100], " grid", d.get("Grid Size", ("", ""))[1], " block", d.get("Block Size", ("", ""))[1])
        for k in KEYS:
            if k in d and d[k][1] != "":
                print(f"  {k:78s} {d[k][1]:>18s} {d[k][0]}")
        print()
    if src:
        rows = list(csv.reader(open(src)))
        for i, r in enumerate(rows):
            if r and r[0] == "Kernel Name":
                name, hdr, start = r[1], rows[i + 1], i + 2
                break
        end = next((j for j in range(start, len(rows)) if rows[j] and rows[j][0] == "Kernel Name"), len(rows))
        body = rows[start:end]
        cols = [j for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = {hdr[j]: sum(int(r[j]) for r in body if len(r) > j and r[j].strip().isdigit()) for j in cols}
        s = sum(tot.values()) or 1
        print("warp-state samples (all), first kernel:", name[:80])
        for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
            print(f"  {k:28s} {100 * v / s:5.1f} %")
        iS, iE = hdr.index("Source"), hdr.index("Instructions Executed")
        iW = [j for j, h in enumerate(hdr) if h.startswith("Warp Stall Sampling (All")][0]
        data = [(int(r[iW]), r[iE], r[iS][:90]) for r in body if len(r) > iW and r[iW].strip().isdigit()]
        t = sum(d[0] for d in data) or 1
        print("hottest instructions by samples:")
        for d in sorted(data, reverse=True)[:10]:
            print(f"  {100 * d[0] / t:5.1f} %  exec {d[1]:>9s}  {d[2]}")
        import collections
        import re
        c = collections.Counter()
        n = 0
        for r in body:
            if len(r) > iE and r[iE].strip().isdigit():
                tks = r[iS].split()
                if not tks:
                    continue
                op = tks[1] if tks[0].startswith("@") and len(tks) > 1 else tks[0]
                c[op.split(".")[0]] += int(r[iE]); n += int(r[iE])
        print(f"dynamic warp instructions {n}; mix:", ", ".join(f"{op} {100 * v / n:.1f}%" for op, v in c.most_common(14)))


if __name__ == "__main__":
    main(sys.argv[1], (sys.argv[2] or None) if len(sys.argv) > 2 else None, int(sys.argv[3]) if len(sys.argv) > 3 else 4)
