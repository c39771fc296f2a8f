#!/usr/bin/env python
"""Top stalled SASS instructions and opcode mix from `ncu -i rep --page source --csv` output.

usage: ncu -i x.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv; python tools/ncu_source_top.py src.csv [N]
"""
import collections
import csv
import sys


def main(path, n=30):
    rows = list(csv.reader(open(path)))
    blocks, cur, hdr = [], None, None
    for r in rows:
        if "Source" in r and "Address" in r:
            hdr = r
            cur = []
            blocks.append(cur)
            continue
        if cur is not None and hdr and len(r) == len(hdr):
            cur.append(r)
    idx = {h: i for i, h in enumerate(hdr)}

    def I(r, k):
        try:
            return int(float(r[idx[k]] or 0))
        except ValueError:
            return 0

    blk = blocks[0]
    tot_s = sum(I(r, "# Samples") for r in blk)
    tot_i = sum(I(r, "Instructions Executed") for r in blk)
    print(f"{len(blocks)} kernel(s) in file; first: {len(blk)} SASS lines, {tot_s} samples, {tot_i} warp-instructions")
    for r in sorted(blk, key=lambda r: -I(r, "# Samples"))[:n]:
        print(f"{I(r, '# Samples'):6d} {100.0 * I(r, '# Samples') / max(tot_s, 1):5.1f}% {I(r, 'Instructions Executed'):9d}  {r[1][:110]}")
    h = collections.Counter()
    for r in blk:
        toks = r[1].split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
        h[op.split(".")[0]] += I(r, "Instructions Executed")
    print("opcode mix (warp-instructions):", ", ".join(f"{k}:{v}" for k, v in h.most_common(18)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
