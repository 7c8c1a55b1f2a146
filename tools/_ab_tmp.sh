timeout 600 python -m pytest tests/test_gpu_light_lists.py -m gpu -q -x 2>&1 | tail -15
for v in "" _v2 _v4u1; do
  export MCPT_LIB=$PWD/monte_carlo_path_tracing_b200/libmcpt$v.so
  timeout 300 python bench.py --steps 3 --warmup 3 --workload veach-mis --no-cpu-baseline --no-extras > gpurun_out/ab$v.json 2> gpurun_out/ab$v.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/ab$v.json").read().strip().splitlines()[-1])
r=j["roofline"]; pl=(r.get("hbm",r))["pipeline"]
print("variant '$v' value %.1f ms/step %.2f ms_shade %.1f ms_trace %.1f"%(j["value"],j["ms_per_step"],pl["ms_shade"],pl["ms_trace"]), r["bound"], "%.3f"%r["frac"])
PY
done
unset MCPT_LIB
bash tools/gpu_launches.sh r2i_veach
