"""Static SASS instruction counts per kernel of libb2sensor.so → profiles/r2_sass_excerpt.txt

    python tools/sass_excerpt.py [path/to/libb2sensor.so]
"""
import collections
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OPS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "SYNCS", "FFMA2", "LDCU", "MUFU", "HMMA", "DMUL", "STL", "LDL"]
HEAD = """# SASS evidence for libb2sensor.so at HEAD (cuobjdump -sass, sm_100a): static instruction counts per kernel.
# UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA engine),
# SYNCS = mbarrier, FFMA2 = packed FP32 FMA, LDCU = uniform constant load, HMMA = warp-level mma.sync (only the optional decode_kernel_wm variant, §9 of the notes),
# STL / LDL = local-memory traffic.
"""


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else glob.glob(os.path.join(ROOT, "*_b200", "libb2sensor.so"))[0]
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    names = {}
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for o in OPS:
                if op == o or (o == "HMMA" and op.startswith("HMMA")):
                    counts[cur][o] += 1
    dem = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
    for k, d in zip(counts, dem):
        names[k] = d
    out = [HEAD]
    for k in sorted(counts, key=lambda k: names[k]):
        c = counts[k]
        out.append(names[k])
        out.append("    instructions=%d  " % c["_total"] + "  ".join(f"{o}={c[o]}" for o in OPS if c[o]))
    path = os.path.join(ROOT, "profiles", "r2_sass_excerpt.txt")
    open(path, "w").write("\n".join(out) + "\n")
    print(path, len(counts), "kernels")


if __name__ == "__main__":
    main()
