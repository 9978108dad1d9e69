#!/usr/bin/env python
"""Reads .ncu-rep captures (ncu --set full) here, without a GPU, and prints the metrics the roofline discussion uses.
    python tools/ncu_summarize.py gpurun_out/r2prof/*.ncu-rep > profiles/r02_ncu_full_summary.txt"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum"]


def main():
    for path in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        if len(rows) < 3:
            print(f"## {path}: unreadable")
            continue
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            rec = dict(zip(hdr, r))
            print(f"## {path}: {rec.get('Kernel Name', '?')}  grid {rec.get('Grid Size', rec.get('launch__grid_size'))} block {rec.get('Block Size', '')}")
            for k in KEYS:
                if k in rec:
                    u = units[hdr.index(k)]
                    print(f"   {k:78s} {rec[k]:>16s} {u}")
        print()


if __name__ == "__main__":
    main()
