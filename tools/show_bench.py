"""Developer tool: print the headline figures of a bench.py JSON line.  usage: python tools/show_bench.py file.json"""
import json
import sys

d = json.loads(open(sys.argv[1]).readline())
print("n_gpus %d  value %.1f  e2e %.1f  e2e_points %.1f  e2e_dense %.1f  ms/step %.2f" % (
    d["n_gpus"], d["value"], d["e2e"]["value"], d["e2e_points"]["value"], d["e2e_dense"]["value"], d["ms_per_step"]))
print("  host_issue", {k: round(v, 1) for k, v in d["host_issue_ms_per_step"].items()}, " seam", d.get("seam_parity"))
for r in d.get("per_rank", []):
    ck = (r.get("clocks") or {}).get("value") or {}
    print("  rank %d ms/step %s  host %s  sm_mhz %s (min %s) power %s reasons %s" % (
        r["rank"], r["ms_per_step"], r["host_issue_ms_per_step"].get("value"), ck.get("sm_mhz"), ck.get("sm_mhz_min"), ck.get("power_w_max"), ck.get("reasons")))
    if r.get("timeline"):
        print("     timeline device", r["timeline"]["device_ms"], " host", r["timeline"]["host_issue_ms"])
if d.get("roofline", {}).get("issue"):
    print("  issue roofline", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in d["roofline"]["issue"].items() if k != "source"})
if "e2e_class" in d:
    print("  e2e_class", {k: v for k, v in d["e2e_class"].items() if k != "what"})
