# usage: bash tools/ab_trace.sh <tag>  -- A/B of the trace-kernel variants: 4-wide vs compressed 8-wide BVH (MCPT_BVH8), and
# the I2F vs PRMT+FADD byte decoding (libmcpt_magic.so, built with -DMCPT_BYTE_MAGIC=1), on the three trace-bound workloads
tag=${1:-ab}
for w in soup-1m interior cornell-box veach-mis; do
  for v in bvh4 bvh8 bvh8magic; do
    case $v in
      bvh4) export MCPT_BVH8=0; unset MCPT_LIB;;
      bvh8) export MCPT_BVH8=1; unset MCPT_LIB;;
      bvh8magic) export MCPT_BVH8=1; export MCPT_LIB=$PWD/monte_carlo_path_tracing_b200/libmcpt_magic.so;;
    esac
    timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_${w}_${v}.json 2> gpurun_out/${tag}_${w}_${v}.err
    python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/${tag}_${w}_${v}.json").read().strip().splitlines()[-1])
    r=j["roofline"]; pl=(r.get("hbm",r))["pipeline"]
    print("$w $v value %.1f Msamples/s  Mrays/s %.0f  ms_trace %.1f ms_shade %.1f"%(j["value"], j["extra"]["Mrays_per_s"], pl["ms_trace"], pl["ms_shade"]))
except Exception as e: print("$w $v failed", e)
PY
  done
done
unset MCPT_BVH8 MCPT_LIB
