"""Group the SASS of one captured launch into regions of equal execution count (loop bodies) and print each
region's share of the executed warp instructions and of the stall samples.
Usage: python scripts/ncu_regions.py gpurun_out/x.ncu-rep"""
import csv
import io
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[1], rows[2:]
iS, iE, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
grp = []
for k, r in enumerate(data):
    e, sm = int(r[iE]), int(r[iSm])
    if grp and abs(e - grp[-1][2]) <= 0.03 * max(e, grp[-1][2], 1):
        g = grp[-1]; g[1] = k; g[3] += e; g[4] += sm
    else:
        grp.append([k, k, e, e, sm])
tot, ts = sum(g[3] for g in grp), sum(g[4] for g in grp)
print("total warp instructions", tot)
for g in grp:
    if g[3] / tot > 0.004:
        print(f"{g[0]:4d}-{g[1]:4d} n={g[1]-g[0]+1:3d} exec/inst={g[2]/1e6:6.2f}M share={100*g[3]/tot:5.1f}% samples={100*g[4]/ts:5.1f}%  {data[g[0]][iS].strip()[:40]}")
