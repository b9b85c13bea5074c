import json,sys
d=json.loads(sys.stdin.readline()); print("value %.1f e2e %.1f" % (d["value"], d["e2e"]["value"]), {k: round(v,1) for k,v in d["kernels_standalone_us"].items()})
