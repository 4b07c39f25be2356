#!/usr/bin/env python
"""Top warp-stall sample sites of each kernel in an `ncu --page source --csv` export.

    ncu -i X.ncu-rep --page source --csv > /tmp/src.csv && python tools/ncu_hotspots.py /tmp/src.csv [N]
"""
import csv
import sys


def main(path, n=40):
    rows = list(csv.reader(open(path)))
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = dict(name=r[1], hdr=None, data=[])
            blocks.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and len(r) == len(cur["hdr"]):
            cur["data"].append(r)
    for b in blocks:
        hdr, data = b["hdr"], b["data"]
        ci = {h: i for i, h in enumerate(hdr)}
        samp, src, ex = ci["# Samples"], ci["Source"], ci["Instructions Executed"]
        stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[samp]) for r in data)
        print(f"=== {b['name'][:100]}: {len(data)} SASS instructions, {tot} samples")
        agg = {s: sum(int(r[ci[s]]) for r in data) for s in stalls}
        print("    stall mix:", ", ".join(f"{k[6:]} {100 * v / max(tot, 1):.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
        top = sorted(range(len(data)), key=lambda i: -int(data[i][samp]))[:n]
        for i in sorted(top):
            r = data[i]
            st = sorted(((int(r[ci[s]]), s[6:]) for s in stalls), reverse=True)[:2]
            print(f"    [{i:5d}] {int(r[samp]):6d} samples  exec {r[ex]:>8s}  {r[src].strip()[:72]:72s} {st}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
