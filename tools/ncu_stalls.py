"""Warp-stall sample totals of one kernel in an ncu report (source page, summed over all SASS rows).
usage: python tools/ncu_stalls.py report.ncu-rep kernel_regex"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1:3]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; tot = {}
for r in rows:
    if r and r[0] in ("Address", "Line No", "#"):
        hdr = r
    elif hdr and len(r) == len(hdr):
        for h, v in zip(hdr, r):
            if h.startswith("stall_") and "Not Issued" not in h:
                try: tot[h] = tot.get(h, 0) + int(v or 0)
                except ValueError: pass
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:12]:
    print(f"{100*v/max(s,1):5.1f}%  {k}")
