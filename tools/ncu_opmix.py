"""Executed warp-instructions of an .ncu-rep by SASS opcode (source page), per record-lane.  Usage: ncu_opmix.py rep n_records"""
import collections, csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw[1:])); hdr = rows[0]
iE, iSrc = hdr.index("Instructions Executed"), hdr.index("Source")
mix = collections.Counter(); tot = 0
for r in rows[1:]:
    try: e = int(r[iE])
    except ValueError: continue
    toks = r[iSrc].strip().split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    op = op.split('.')[0]
    mix[op] += e; tot += e
nrec = float(sys.argv[2])
print("total warp-inst", tot, " per 32 records", round(tot / (nrec / 32), 1))
for op, e in mix.most_common(45):
    print(f"{op:12s} {e:10d} {100*e/tot:5.1f}%  {e/(nrec/32):7.1f} /rec-lane")
