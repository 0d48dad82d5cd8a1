"""Executed warp-instructions and stall samples per CUDA source line of one kernel, in line order (duplicates from
several launches / inlined copies merged).  usage: python tools/ncu_lines.py report.ncu-rep kernel_regex file lo hi"""
import csv, io, subprocess, sys
rep, kre, fn, lo, hi = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, fname = None, ""
acc, ti, ts = {}, 0, 0
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
        if fname == fn and lo <= int(r[0]) <= hi:
            a = acc.setdefault(int(r[0]), [0, 0, r[1].strip()[:90]]); a[0] += i_; a[1] += s_
for l in sorted(acc):
    i_, s_, src = acc[l]
    if i_ or s_:
        print(f"{l:4d} {100*i_/ti:5.2f}% inst {100*s_/ts:5.2f}% smp  {src}")
