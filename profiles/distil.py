"""Copy what is worth keeping from a capture (gpurun_out/<tag>, scratch) into profiles/<tag> (tracked):
bench lines, the ncu launch list, per-kernel `--page details` text, the handful of raw metrics the
roofline quotes, and summary.json -- which bench.py reads for `roofline.traffic`.

    python profiles/distil.py <tag> [--current]      (--current: also write profiles/current.json)
"""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}


def raw_metrics(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"]
    if not hdr:
        return {}
    h = hdr[0]
    return {n: (u, v) for n, u, v in zip(rows[h], rows[h + 1], rows[h + 2])}


def main():
    tag = sys.argv[1]
    src, dst = os.path.join(ROOT, "gpurun_out", tag), os.path.join(ROOT, "profiles", tag)
    os.makedirs(dst, exist_ok=True)
    for f in ("bench.json", "bench_reference.json", "launches.csv", "pytest.log"):
        if os.path.exists(os.path.join(src, f)):
            shutil.copy(os.path.join(src, f), os.path.join(dst, f))
    summary = {}
    for f in sorted(os.listdir(src)):
        if f.endswith("_details.txt"):
            shutil.copy(os.path.join(src, f), os.path.join(dst, f))
        if f.endswith("_raw.csv"):
            k = f[:-len("_raw.csv")]
            m = raw_metrics(os.path.join(src, f))
            with open(os.path.join(dst, k + "_raw_metrics.txt"), "w") as out:
                for name in KEEP:
                    if name in m:
                        out.write(f"{name}\t{m[name][1]}\t{m[name][0]}\n")

            def val(name):
                if name not in m:
                    return None
                u, v = m[name]
                return float(v.replace(",", "")) * UNIT.get(u, 1.0)

            rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
            summary[k] = {"duration_s": val("gpu__time_duration.sum"), "dram_bytes_read": rd, "dram_bytes_write": wr,
                          "dram_bytes": (rd or 0) + (wr or 0) if rd is not None or wr is not None else None,
                          "warp_instructions": val("smsp__inst_executed.sum")}
    json.dump(summary, open(os.path.join(dst, "summary.json"), "w"), indent=1)
    if "--current" in sys.argv:
        json.dump({"tag": tag, "kernels": summary}, open(os.path.join(ROOT, "profiles", "current.json"), "w"), indent=1)
    print(json.dumps(summary, indent=1))


if __name__ == "__main__":
    main()
