#!/usr/bin/env python
"""Achieved DRAM / L2 traffic of the HBM-bound kernels from `ncu --set full` reports (raw-page CSVs): one row per launch.

    for f in gpurun_out/r2f_hbm_*.ncu-rep; do ncu -i $f --page raw --csv > /tmp/$(basename $f).csv; done
    python tools/ncu_hbm_table.py /tmp/r2f_hbm_*.csv > profiles/r2_ncu_hbm_kernels.md
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(paths):
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.isfile(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    print("| kernel | grid x block | time us | dram read MB | dram write MB | dram GB/s (of %.0f measured) | L2 read+write MB | L2 GB/s | "
          "`dram__throughput` %% | `lts__throughput` %% | warps active %% |" % pk)
    print("|---|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
    for p in paths:
        rows = list(csv.reader(open(p)))
        hdr, data = rows[0], rows[2:]
        col = {h: i for i, h in enumerate(hdr)}
        units = rows[1]

        def val(r, k, scale=None):
            if k not in col:
                return float("nan")
            v = float(r[col[k]].replace(",", "")) if r[col[k]] not in ("", "n/a") else float("nan")
            u = units[col[k]].lower()
            if scale == "bytes":
                v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            if scale == "us":
                v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1)
            return v
        for r in data:
            name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
            t = val(r, "gpu__time_duration.sum", "us")
            rd, wr = val(r, "dram__bytes_read.sum", "bytes"), val(r, "dram__bytes_write.sum", "bytes")
            l2 = (val(r, "lts__t_sectors_srcunit_tex_op_read.sum") + val(r, "lts__t_sectors_srcunit_tex_op_write.sum")) * 32
            grid = f'{r[col["launch__grid_size"]]} x {r[col["launch__block_size"]]}'
            print(f"| `{name}` | {grid} | {t:.2f} | {rd / 1e6:.2f} | {wr / 1e6:.2f} | {(rd + wr) / t / 1e3:.0f} ({(rd + wr) / t / 1e3 / pk * 100:.0f} %) | "
                  f"{l2 / 1e6:.2f} | {l2 / t / 1e3:.0f} | {val(r, 'FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
                  f"{val(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
                  f"{val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} |")


if __name__ == "__main__":
    main(sys.argv[1:])
