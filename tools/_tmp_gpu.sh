timeout 600 python -m pytest tests -m gpu -x -q -k "light_tree" 2>&1 | tail -5
for w in veach-mis interior; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload $w --light-tree > gpurun_out/lt1_$w.json 2>gpurun_out/lt1_$w.err
python - <<PY
import json
for t in ("lt1",):
    try:
        j=json.load(open("gpurun_out/%s_$w.json"%t)); p=j["roofline"]["pipeline"]; print(t, "$w", round(j["value"],1), "e2e", round(j["e2e"]["value"],1), "ms/step", round(j["ms_per_step"],2), "shade", round(p["ms_shade"]/3,2), "trace", round(p["ms_trace"]/3,2), "launches", j["gpu_launches"])
    except Exception as e: print(t, "$w failed", e, open("gpurun_out/%s_$w.err"%t).read()[-800:])
PY
done
