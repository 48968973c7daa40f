"""Flatten an `ncu --set full` report into the per-launch text summary kept under profiles/.

    ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/x.csv && python tools/ncu_summary.py /tmp/x.csv > profiles/...
"""
import csv
import sys

WANT = ["Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print(f"{'Kernel Name':72s} {r[hdr.index('Kernel Name')][:110]}")
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:72s} {r[i]} {units[i]}")
    print()
