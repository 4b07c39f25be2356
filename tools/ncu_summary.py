#!/usr/bin/env python
"""Key metrics of an `ncu --set full` report as a markdown table (one column per captured launch).

    ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv && python tools/ncu_summary.py /tmp/raw.csv > profiles/rNN_X.md
"""
import csv
import sys

WANT = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__cycles_active.avg",
]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    names = [r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("<unnamed>::", "") for r in data]
    print("| metric | unit | " + " | ".join(f"`{n}` #{i}" for i, n in enumerate(names)) + " |")
    print("|---|---|" + "---:|" * len(data))
    for w in WANT:
        if w not in col:
            continue
        i = col[w]
        print(f"| `{w}` | {units[i]} | " + " | ".join(r[i] for r in data) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
