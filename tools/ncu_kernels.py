"""One line per profiled kernel launch: duration, instructions, issue/pipe utilisation, occupancy, DRAM bytes.
usage: python tools/ncu_kernels.py report.ncu-rep"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out))); h = r[0]; idx = {n: i for i, n in enumerate(h)}
cols = [("us", "gpu__time_duration.sum"), ("Minst", "smsp__inst_executed.sum"),
        ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        ("fmaH%", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        ("fma%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        ("lsu%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        ("xu%", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("rdMB", "dram__bytes_read.sum"), ("wrMB", "dram__bytes_write.sum"), ("regs", "launch__registers_per_thread"),
        ("l1hit%", "l1tex__t_sector_hit_rate.pct"), ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")]
print("kernel".ljust(22) + "".join(n.rjust(9) for n, _ in cols))
for row in r[2:]:
    name = row[idx["Kernel Name"]].split("(")[0][:21]
    vals = []
    for n, m in cols:
        v = row[idx[m]] if m in idx else ""
        try:
            f = float(v.replace(",", ""))
            if n == "Minst": f /= 1e6
            vals.append(f"{f:9.2f}")
        except ValueError:
            vals.append(v[:8].rjust(9))
    print(name.ljust(22) + "".join(vals))
