"""Developer tool: the few numbers of a bench line one looks at between kernel iterations.  usage: show_bench.py line.json"""
import json, sys
d = json.loads(open(sys.argv[1]).readline()) if len(sys.argv) > 1 else json.loads(sys.stdin.readline())
print("value %.1f  e2e %.1f  e2e_points %.1f  e2e_dense %.1f  ms/step %.2f" % (
    d["value"], d["e2e"]["value"], d.get("e2e_points", {}).get("value", 0), d.get("e2e_dense", {}).get("value", 0), d["ms_per_step"]))
print("standalone us", {k: round(v, 1) for k, v in d["kernels_standalone_us"].items()}, "sum %.1f" % sum(d["kernels_standalone_us"].values()))
print("pipelined  us", {k: round(v, 1) for k, v in d["kernels_avg_us"].items()})
print("roofline frac %.4f whole %.4f  launches/kf %.2f  host issue %s" % (
    d["roofline"]["frac"], d["roofline"]["whole_path_frac"], d.get("launches_per_keyframe", 0),
    {k: round(v, 1) for k, v in d.get("host_issue_ms_per_step", {}).items()}))
if "seam_parity" in d:
    print("seam", d["seam_parity"])
