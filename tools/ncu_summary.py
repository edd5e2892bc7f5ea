"""Summarise an .ncu-rep (read here, no GPU): python tools/ncu_summary.py rep.ncu-rep [metric-regex]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else
                 r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                 r"sm__throughput\.avg\.pct|sm__warps_active\.avg\.pct_of_peak_sustained_active|launch__registers_per_thread|"
                 r"launch__shared_mem_per_block_dynamic|sm__pipe_tensor.*cycles_active.*pct|lts__t_sector_hit_rate\.pct|"
                 r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|smsp__inst_executed\.sum|sm__inst_executed_pipe_(fma|lsu|alu).*\.sum$|"
                 r"smsp__average_warps_issue_stalled_.*_per_issue_active|lts__t_bytes\.sum|sm__cycles_elapsed\.max)")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:90], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for i, h in enumerate(hdr):
        if pat.search(h) and r[i] not in ("", "0"):
            print(f"   {h:90s} {r[i]:>16s} {units[i]}")
