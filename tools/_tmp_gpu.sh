timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in veach-mis; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/hz_$w.json 2>gpurun_out/hz_$w.err
python - <<PY
import json
try:
    j=json.load(open("gpurun_out/hz_$w.json")); p=j["roofline"]["pipeline"]; print("$w", round(j["value"],1), "e2e", round(j["e2e"]["value"],1), "ms/step", round(j["ms_per_step"],2), "shade", round(p["ms_shade"]/3,2), "trace", round(p["ms_trace"]/3,2), "launches", j["gpu_launches"])
except Exception as e: print("$w failed", e, open("gpurun_out/hz_$w.err").read()[-800:])
PY
done
