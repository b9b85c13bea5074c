"""Developer tool (CPU): summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,
dram__bytes_read.sum,dram__bytes_write.sum --csv --log-file X.csv python bench.py ...`) per kernel, and derive what
bench.py reads from profiles/k1_traffic.json: DRAM bytes per launch of the K1 stage and warp instructions per key frame.

    python tools/ncu_launch_summary.py profiles/r02c_launches.csv [--write-traffic COMMIT]

Per-launch times in such a list are cold-cache and serialised: the kernels' SHARES of a key frame are what must agree
with bench.py's `kernels_standalone_share`, not the absolute durations.
"""
import csv
import json
import os
import re
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

STAGE = [("k0_gradient", "grad"), ("k1_search", "search"), ("k1_fuse", "fuse"), ("k23_intra", "intra"),
         ("k4_inter_check", "inter"), ("k5_pointset", "points"), ("k7_pack", "pack")]


def to_float(v):
    return float(v.replace(",", ""))


def unit_scale(metric, unit):
    u = unit.strip().lower()
    if metric.startswith("gpu__time_duration"):
        return {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}.get(u, 1.0)  # -> us
    if metric.startswith("dram__bytes"):
        return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)  # -> bytes
    return 1.0


def load(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    launches = defaultdict(dict)
    names = {}
    for r in rd:
        try:
            i = int(r["ID"])
        except (KeyError, ValueError):
            continue
        names[i] = r["Kernel Name"]
        m = r["Metric Name"]
        launches[i][m] = to_float(r["Metric Value"]) * unit_scale(m, r.get("Metric Unit", ""))
    for i in sorted(launches):
        rows.append((names[i], launches[i]))
    return rows


def main():
    path = sys.argv[1]
    rows = load(path)
    per = defaultdict(lambda: defaultdict(list))
    for name, m in rows:
        short = re.sub(r"\(.*", "", name).replace("sdpm::", "")
        short = re.sub(r"^void ", "", short)
        for k, v in m.items():
            per[short][k].append(v)
    out = {}
    print(f"{len(rows)} launches in {path}")
    print(f"{'kernel':34s} {'n':>4s} {'avg us':>9s} {'warp inst':>12s} {'dram rd MB':>11s} {'dram wr MB':>11s}")
    for k in sorted(per, key=lambda k: -sum(per[k].get("gpu__time_duration.sum", [0]))):
        d = per[k]
        n = len(d.get("gpu__time_duration.sum", []))
        avg = lambda key: (sum(d[key]) / len(d[key])) if d.get(key) else 0.0
        out[k] = {"launches": n, "avg_us": avg("gpu__time_duration.sum"), "warp_instructions": avg("smsp__inst_executed.sum"),
                  "dram_read_bytes": avg("dram__bytes_read.sum"), "dram_write_bytes": avg("dram__bytes_write.sum")}
        print(f"{k[:34]:34s} {n:4d} {out[k]['avg_us']:9.1f} {out[k]['warp_instructions']:12.0f} {out[k]['dram_read_bytes']/1e6:11.2f} {out[k]['dram_write_bytes']/1e6:11.2f}")
    # per key frame: one launch of each stage kernel
    stage = {}
    for prefix, label in STAGE:
        ks = [k for k in out if k.startswith(prefix)]
        if ks:
            k = max(ks, key=lambda k: out[k]["launches"])
            stage[label] = out[k]
    path_stages = [s for s in ("grad", "search", "fuse", "intra", "inter", "points") if s in stage]
    tot_us = sum(stage[s]["avg_us"] for s in path_stages)
    tot_inst = sum(stage[s]["warp_instructions"] for s in path_stages)
    print("per key frame (one launch per stage): %.1f us, %.1f M warp instructions" % (tot_us, tot_inst / 1e6))
    print("shares of the serialised time:", {s: round(stage[s]["avg_us"] / tot_us, 3) for s in path_stages})
    summary = {"source": os.path.relpath(path, ROOT), "per_kernel": out,
               "per_keyframe": {"us_serialised": tot_us, "warp_instructions": tot_inst,
                                "share": {s: stage[s]["avg_us"] / tot_us for s in path_stages}}}
    with open(os.path.splitext(path)[0] + "_summary.json", "w") as f:
        json.dump(summary, f, indent=1)
    if "--write-traffic" in sys.argv:
        commit = sys.argv[sys.argv.index("--write-traffic") + 1]
        k1 = stage["search"], stage["fuse"]
        traffic = sum(s["dram_read_bytes"] + s["dram_write_bytes"] for s in k1)
        rel = os.path.relpath(path, ROOT)
        d = {"dram_bytes_per_launch": int(traffic),
             "parts": {"k1_search": int(k1[0]["dram_read_bytes"] + k1[0]["dram_write_bytes"]),
                       "k1_fuse": int(k1[1]["dram_read_bytes"] + k1[1]["dram_write_bytes"])},
             "warp_instructions_per_keyframe": int(tot_inst),
             "warp_instructions_parts": {s: int(stage[s]["warp_instructions"]) for s in path_stages},
             "warp_instructions_source": f"ncu smsp__inst_executed.sum, one launch of each stage kernel (720p, N=10), averaged over the launch list {rel}",
             "source": f"ncu --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum per launch of k1_search and k1_fuse<10> (720p, N=10), averaged over the launch list {rel}",
             "algorithmic_bytes_per_launch": 98611200, "camera": "720p", "neighbours": 10,
             "commit": commit}
        with open(os.path.join(ROOT, "profiles", "k1_traffic.json"), "w") as f:
            json.dump(d, f, indent=1)
        print("wrote profiles/k1_traffic.json:", d)


if __name__ == "__main__":
    main()
