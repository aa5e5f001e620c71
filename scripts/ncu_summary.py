#!/usr/bin/env python
"""Summarises an ncu report (gpurun_out/*.ncu-rep, `--set full --import-source on`) into the small CSV that is
committed under profiles/: the headline metrics of every profiled launch, its warp-stall breakdown (sampling) and
the ten SASS lines with most samples. Usage: python scripts/ncu_summary.py <report.ncu-rep> <out.csv>"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__thread_inst_executed_per_inst_executed.ratio"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    w = csv.writer(open(out, "w", newline=""))
    w.writerow(["report", rep.split("/")[-1]])
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        w.writerow(["kernel", name])
        for i, h in enumerate(hdr):
            if h in WANT:
                w.writerow(["metric", h, units[i], r[i]])
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    if len(rows) > 2:
        hdr = rows[1]
        idx = {h: i for i, h in enumerate(hdr)}
        data = [r for r in rows[2:] if len(r) == len(hdr)]
        stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[idx["# Samples"]]) for r in data) or 1
        w.writerow(["source_page", "first profiled kernel", "samples", tot, "instructions executed", sum(int(r[idx["Instructions Executed"]]) for r in data)])
        for h in sorted(stalls, key=lambda h: -sum(int(r[idx[h]]) for r in data)):
            v = sum(int(r[idx[h]]) for r in data)
            if v:
                w.writerow(["stall", h, v, "%.1f%%" % (100.0 * v / tot)])
        base = int(data[0][idx["Address"]], 16)
        for r in sorted(data, key=lambda r: -int(r[idx["# Samples"]]))[:10]:
            w.writerow(["hot_sass", hex(int(r[idx["Address"]], 16) - base), r[idx["Source"]].strip(), r[idx["# Samples"]], r[idx["Instructions Executed"]]])
    print("wrote", out)


if __name__ == "__main__":
    main()
