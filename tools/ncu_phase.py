"""Executed warp-instructions / stall samples of one kernel grouped by source-line ranges.
usage: python tools/ncu_phase.py report.ncu-rep kernel_regex file.inl name:lo-hi [name:lo-hi ...]"""
import csv, io, subprocess, sys
rep, kre, fn = sys.argv[1:4]
ranges = []
for a in sys.argv[4:]:
    n, r = a.split(":"); lo, hi = r.split("-"); ranges.append((n, int(lo), int(hi)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, fname = None, ""
tot = {}; ti = ts = 0
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r; ie, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
    elif hdr and r and r[0].isdigit():
        try:
            i_, s_ = int(r[ie] or 0), int(r[sm] or 0)
        except ValueError:
            continue
        ti += i_; ts += s_
        key = "other:" + fname
        if fname == fn:
            for n, lo, hi in ranges:
                if lo <= int(r[0]) <= hi:
                    key = n; break
        a = tot.setdefault(key, [0, 0]); a[0] += i_; a[1] += s_
print("total warp-inst", ti, "samples", ts)
for k, (i_, s_) in sorted(tot.items(), key=lambda x: -x[1][0]):
    print(f"{100*i_/max(ti,1):5.1f}% inst {100*s_/max(ts,1):5.1f}% smp  {k}")
