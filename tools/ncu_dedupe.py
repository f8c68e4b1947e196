"""Reduce the per-launch summary of tools/ncu_summary.py to one entry per kernel of the library.

    python tools/ncu_dedupe.py <summary.json> <out.json>

tools/prof_all.py launches some kernels several times (LSTM forward steps, the bitonic merge passes); the first launch of each
(kernel, grid size) pair is kept.  One-CTA launches of decode_kernel are the oversize-retry launch over an empty list; the two
launches of decode_kernel<18,1,1,1> are Avro-A and Avro-B in prof_all's order.
"""
import json
import sys


def main(src, dst):
    rows = json.load(open(src))
    seen, out, avro = set(), [], 0
    for e in rows:
        grid = e.get("launch__grid_size", ["", ""])[0]
        key = (e["kernel"], grid)
        name = e["kernel"]
        if (name.startswith("void decode_kernel<18, 1, 1, 1>") or name.startswith("void decode_kernel_zs<18, 1, 1>")) and grid != "1":
            avro += 1
            key = (name, grid, avro)
            e["note"] = "Avro-A, 2^20 records" if avro == 1 else "Avro-B, 2^20 records"
        if key in seen:
            continue
        seen.add(key)
        if "decode_kernel" in name and grid == "1":
            e["note"] = "oversize-retry launch, list empty (one CTA)"
        out.append(e)
    json.dump(out, open(dst, "w"), indent=1)
    print(f"{len(rows)} launches -> {len(out)} kernels")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
