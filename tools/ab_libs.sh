# usage: bash tools/ab_libs.sh <workload> <suffix> [<suffix> ...] -- bench one workload with libmcpt<suffix>.so variants ("-" = the shipped library)
w=$1; shift
for v in "$@"; do
  [ "$v" = "-" ] && v=""
  export MCPT_LIB=$PWD/monte_carlo_path_tracing_b200/libmcpt$v.so
  timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline --no-extras > gpurun_out/ab_$w$v.json 2> gpurun_out/ab_$w$v.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_$w$v.json").read().strip().splitlines()[-1])
    r=j["roofline"]; pl=(r.get("hbm",r))["pipeline"]
    print("$w variant '$v' value %.1f ms/step %.2f ms_shade %.1f ms_trace %.1f"%(j["value"],j["ms_per_step"],pl["ms_shade"],pl["ms_trace"]), r["bound"], "%.3f"%r["frac"])
except Exception as e: print("$w '$v' failed", e); print(open("gpurun_out/ab_$w$v.err").read()[-800:])
PY
done
unset MCPT_LIB
