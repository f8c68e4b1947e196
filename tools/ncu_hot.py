"""Top SASS instructions of an .ncu-rep by stall samples, with executed counts. Usage: ncu_hot.py rep [N]"""
import csv, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw[1:]))
hdr = rows[0]
iS, iN, iE, iSrc = hdr.index("# Samples"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed"), hdr.index("Source")
data = []
tot_s = tot_e = 0
for k, r in enumerate(rows[1:]):
    try:
        s = int(r[iS]); e = int(r[iE])
    except ValueError:
        continue
    tot_s += s; tot_e += e
    data.append((s, e, k, r[iSrc]))
print("total samples", tot_s, "total warp-inst", tot_e, "instructions", len(data))
# cumulative by region: print top N
for s, e, k, src in sorted(data, reverse=True)[:N]:
    print(f"{s:7d} {100*s/tot_s:5.1f}%  exec {e:9d}  #{k:5d}  {src[:110]}")
