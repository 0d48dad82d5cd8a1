"""Per-kernel SASS opcode histogram of liborbx.so (cuobjdump -sass), so that the UTCQMMA / LDTM (tcgen05, tensor memory) / UTMALDG / UBLKCP / SYNCS / IDP /
VIMNMX3 / VABSDIFF4 claims of DESIGN.md can be checked without rebuilding.
    python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "orb_slam_b200/liborbx.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, hist = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); hist[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
print("# per-kernel SASS opcode histogram of orb_slam_b200/liborbx.so (cuobjdump -sass, sm_100a), mnemonic without modifiers;")
print("# top 14 per kernel + the Blackwell / Hopper data-movement and SIMD opcodes whatever their rank")
key = ["UTCQMMA", "UTCHMMA", "LDTM", "UTCBAR", "UTCATOMSWS", "UTMALDG", "UBLKCP", "SYNCS", "LDGSTS", "IDP", "VIMNMX3", "VIMNMX", "VABSDIFF4", "POPC", "REDUX", "ATOMS", "ACQBULK", "PREEXIT"]
for k, c in hist.items():
    name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
    print(f"{name}  [{sum(c.values())} instr]")
    print("    top: " + ", ".join(f"{o} {n}" for o, n in c.most_common(14)))
    print("    key: " + ", ".join(f"{o} {c[o]}" for o in key if c[o]))
