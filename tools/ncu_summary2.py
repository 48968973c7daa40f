"""Per-kernel summary of one or more .ncu-rep files (ncu --page raw --csv): duration, tensor-pipe activity, DRAM bytes,
shared-memory wavefronts / bank conflicts, issue activity.  Also writes the {kernel: {...}} table bench.py reads
(profiles/traffic.json) when given --json PATH.

    python tools/ncu_summary2.py a.ncu-rep [b.ncu-rep ...] [--json profiles/traffic.json]
"""
import csv, io, json, re, subprocess, sys

WANT = [("gpu__time_duration.sum", "us", 1e-3 if False else 1.0),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct", 1.0),
        ("sm__inst_executed_pipe_tensor.sum", None, 1.0),
        ("dram__bytes_read.sum", "dram_read", 1.0), ("dram__bytes_write.sum", "dram_write", 1.0),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct", 1.0),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts", 1.0),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts", 1.0),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct", 1.0),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct", 1.0),
        ("launch__registers_per_thread", "regs", 1.0), ("launch__grid_size", "grid", 1.0),
        ("sm__cycles_active.avg", "sm_cycles", 1.0)]
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "ms": 1e3, "msecond": 1e3, "usecond": 1.0, "nsecond": 1e-3, "second": 1e6}


def short(n):
    n = re.sub(r"^void ", "", n)
    return re.sub(r"\(.*$", "", n)


def rows_of(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = {"kernel": short(r[hdr.index("Kernel Name")])}
        for m, key, _ in WANT:
            if m in hdr and key:
                i = hdr.index(m)
                try:
                    d[key] = float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)
                except ValueError:
                    pass
        yield d


def main(argv):
    js = None
    if "--json" in argv:
        js = argv[argv.index("--json") + 1]
        argv = [a for a in argv if a not in ("--json", js)]
    agg = {}
    for p in argv:
        for d in rows_of(p):
            agg.setdefault(d["kernel"], []).append(d)
    table = {}
    print(f"{'kernel':52s} {'n':>2s} {'us':>8s} {'tensor%':>8s} {'dram MB':>9s} {'dram%':>6s} {'smem wf':>10s} {'conflict%':>9s} {'issue%':>7s} {'regs':>5s}")
    for k, ds in sorted(agg.items(), key=lambda kv: -sum(d.get("us", 0) for d in kv[1])):
        n = len(ds)
        mean = lambda key: sum(d.get(key, 0.0) for d in ds) / n
        dram = mean("dram_read") + mean("dram_write")
        wf = mean("smem_wavefronts")
        print(f"{k[:52]:52s} {n:2d} {mean('us'):8.1f} {mean('tensor_pipe_pct'):8.1f} {dram / 1e6:9.1f} {mean('dram_pct'):6.1f} {wf:10.0f} "
              f"{100 * mean('smem_conflicts') / wf if wf else 0:9.1f} {mean('issue_active_pct'):7.1f} {mean('regs'):5.0f}")
        table[k] = {"dram_bytes_per_launch": dram, "tensor_pipe_active_pct": round(mean("tensor_pipe_pct"), 2), "us_under_ncu": round(mean("us"), 1),
                    "launches_captured": n}
    if js:
        json.dump(table, open(js, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1:])
