export MCPT_DEBUG_TIMING=1
for i in 1 2 3 4; do
timeout 300 python bench.py --cpu-seconds 2 > gpurun_out/alt_cpu_$i.json 2>gpurun_out/alt_cpu_$i.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/alt_nocpu_$i.json 2>gpurun_out/alt_nocpu_$i.err
done
python - <<PY
import json
for i in (1,2,3,4):
  for w in ("cpu","nocpu"):
    j=json.load(open("gpurun_out/alt_%s_%d.json"%(w,i))); print(w, i, round(j["value"],1), j["extra"]["host_ms_per_step"])
    if j["value"] < 1100:
        print("".join(open("gpurun_out/alt_%s_%d.err"%(w,i)).readlines()[3:8]))
PY
