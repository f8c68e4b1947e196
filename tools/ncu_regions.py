"""Executed warp-instructions and stall samples per region of a kernel's SASS (buckets of N instructions)."""
import csv, subprocess, sys
rep = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 100
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw[1:])); hdr = rows[0]
iS, iE, iSrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
ex, sm, first = {}, {}, {}
k = 0
for r in rows[1:]:
    try: s = int(r[iS]); e = int(r[iE])
    except ValueError: continue
    b = k // B; ex[b] = ex.get(b, 0) + e; sm[b] = sm.get(b, 0) + s; first.setdefault(b, r[iSrc][:60]); k += 1
te, ts = sum(ex.values()), sum(sm.values())
for b in sorted(ex):
    if ex[b] * 200 > te or sm[b] * 100 > ts:
        print(f"#{b*B:5d}-{b*B+B-1:5d} exec {100*ex[b]/te:5.1f}%  samples {100*sm[b]/ts:5.1f}%   {first[b]}")
