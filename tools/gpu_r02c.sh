#!/bin/bash
# round-2 evidence run on one B200: GPU suite (verbose), default bench line, prebuilt variants, ncu launch list with
# instruction / DRAM counters, ncu --set full of the search kernel.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
TAG=${TAG:-r02c}
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv,noheader > gpurun_out/${TAG}_gpu.txt 2>&1; nproc >> gpurun_out/${TAG}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -rs -v > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -4 gpurun_out/${TAG}_pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<'P'
import json,os
t=os.environ.get("TAG","r02c")
try:
    d=json.loads(open(f"gpurun_out/{t}_bench.json").readline())
    print("value %.1f e2e %.1f e2e_points %.1f e2e_dense %.1f" % (d["value"], d["e2e"]["value"], d["e2e_points"]["value"], d["e2e_dense"]["value"]))
    print("host_issue", {k: round(v,1) for k,v in d["host_issue_ms_per_step"].items()}, "launches/kf", d["launches_per_keyframe"])
    print("standalone_us", {k: round(v,1) for k,v in d["kernels_standalone_us"].items()})
    print("configs", {k: (round(v["value"]), round(v["e2e"])) for k,v in d.get("configs",{}).items()})
    print("cpu", d.get("cpu_baseline"))
except Exception as e:
    print("bench line unreadable:", e); print(open(f"gpurun_out/{t}_bench.err").read()[-3000:])
P
[ -n "$SKIP_VARIANTS" ] || bash tools/gpu_variants_run.sh 2>&1 | tee gpurun_out/${TAG}_variants.txt
CMD="python bench.py --nkf 24 --steps 2 --warmup 1 --no-cpu-baseline --no-configs"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1; echo "ncu list rc=$?"
for k in ${NCU_KERNELS:-k1_search}; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 30 -c 1 -f -o gpurun_out/${TAG}_$k $CMD > gpurun_out/${TAG}_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
ls -la gpurun_out | head -40
