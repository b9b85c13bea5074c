"""Developer tool (CPU): turn an `ncu --set full` report into the small JSON summary kept under profiles/.

    python tools/ncu_full_summary.py gpurun_out/r02c_k1_search.ncu-rep profiles/r02c_k1_search_ncu_full.json

Reads the report with `ncu -i <rep> --page raw --csv` (works without a GPU) and keeps the counters DESIGN.md argues with.
"""
import csv
import io
import json
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_eligible.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__block_size", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "launch__shared_mem_per_block_static",
    "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
]
KEEP_PREFIX = ["smsp__average_warps_issue_stalled_"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units = rows[hdr], rows[hdr + 1]
    res = []
    for r in rows[hdr + 2:]:
        if len(r) != len(names):
            continue
        d = {"report": rep, "kernel": r[names.index("Kernel Name")]}
        for n, u, v in zip(names, units, r):
            if n in KEEP or any(n.startswith(p) and n.endswith("_per_issue_active.ratio") for p in KEEP_PREFIX):
                key = f"{n} [{u}]" if u else n
                try:
                    d[key] = float(v.replace(",", ""))
                except ValueError:
                    d[key] = v
        res.append(d)
    with open(out, "w") as f:
        json.dump(res[0] if len(res) == 1 else res, f, indent=1)
    for d in res:
        print(d["kernel"][:60], {k: v for k, v in d.items() if k.split(" ")[0] in ("gpu__time_duration.sum", "smsp__inst_executed.sum",
              "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
              "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active")})


if __name__ == "__main__":
    main()
