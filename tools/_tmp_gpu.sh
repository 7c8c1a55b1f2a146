for v in $VARIANTS; do
  cp build_variants/libmcpt_$v.so monte_carlo_path_tracing_b200/libmcpt.so
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - <<PY
import json
j=json.load(open("gpurun_out/var_$v.json"))
p=j["roofline"].get("pipeline",{})
print("$v", "value %.1f e2e %.1f ms_shade %.2f ms_trace %.2f"%(j["value"], j["e2e"]["value"], p.get("ms_shade",0)/3, p.get("ms_trace",0)/3))
PY
done
cp build_variants/libmcpt_base.so monte_carlo_path_tracing_b200/libmcpt.so
