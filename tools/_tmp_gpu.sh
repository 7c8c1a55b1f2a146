for v in $VARIANTS; do
  cp build_variants/libmcpt_$v.so monte_carlo_path_tracing_b200/libmcpt.so
  for w in soup-1m interior cornell-box veach-mis; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/var_${v}_$w.json 2> gpurun_out/var_$v.err
  python - <<PY
import json
j=json.load(open("gpurun_out/var_${v}_$w.json"))
p=j["roofline"].get("pipeline",{})
print("$v $w", "value %.1f ms_shade %.2f ms_trace %.2f"%(j["value"], p.get("ms_shade",0)/3, p.get("ms_trace",0)/3))
PY
  done
done
cp build_variants/libmcpt_nopf.so monte_carlo_path_tracing_b200/libmcpt.so
