# usage: bash tools/gpu_quick.sh <tag> [pytest-args] -- GPU tests (optionally a subset) + the four bench workloads, no ncu
tag=${1:-q}
shift
timeout 1500 python -m pytest tests -m gpu -q -x "$@" > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${tag}_pytest_gpu.log
for w in veach-mis cornell-box soup-1m interior; do
  timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/${tag}_bench_$w.json").read().strip().splitlines()[-1])
    r=j["roofline"]; pl=(r.get("hbm",r))["pipeline"]
    print("$w value %.1f  ms/step %.2f  e2e %.1f  launches %d  ms_shade %.1f ms_trace %.1f"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j.get("gpu_launches",0),pl["ms_shade"],pl["ms_trace"]), r["bound"], "%.3f"%r["frac"], j["extra"]["bvh"].get("light_lists"), j.get("fp32",{}).get("evaluations_per_sweep"))
except Exception as e: print("$w failed", e); print(open("gpurun_out/${tag}_bench_$w.err").read()[-1500:])
PY
done
