"""Per-CUDA-source-line executed instructions / stall samples of one kernel in an ncu report.
usage: python tools/ncu_hot.py report.ncu-rep kernel_regex [top_n]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
items, hdr, fname, done = [], None, "", False
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Function Name":
        pass
    elif r and r[0] == "Line No":
        hdr = r
        ie, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
    elif hdr and r and r[0] not in ("",) and r[0].isdigit():
        try:
            items.append((int(r[ie] or 0), int(r[sm] or 0), fname, r[0], r[1].strip()[:100]))
            done = True
        except ValueError:
            pass
ti = sum(i[0] for i in items); ts = sum(i[1] for i in items)
print("total warp-inst", ti, "samples", ts)
for i_, s_, f, l, src in sorted(items, key=lambda x: -x[0])[:n]:
    print(f"{100*i_/max(ti,1):5.1f}% inst {100*s_/max(ts,1):5.1f}% smp  {f}:{l}: {src}")
