timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in veach-mis interior; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/carry_$w.json 2>/dev/null
python - <<PY
import json
j=json.load(open("gpurun_out/carry_$w.json")); p=j["roofline"]["pipeline"]; print("$w", round(j["value"],1), "shade", round(p["ms_shade"]/3,2), "trace", round(p["ms_trace"]/3,2), "sweeps", j["fp32"]["light_table_sweeps"])
PY
done
