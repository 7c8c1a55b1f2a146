# usage: bash tools/ab_auto.sh <tag> -- what the per-scene layout calibration picks and what it yields, next to both forced layouts
tag=${1:-auto}
for w in soup-1m interior cornell-box veach-mis; do
  for v in auto bvh4 bvh8; do
    case $v in
      auto) unset MCPT_BVH8;;
      bvh4) export MCPT_BVH8=0;;
      bvh8) export MCPT_BVH8=1;;
    esac
    timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_${w}_${v}.json 2> gpurun_out/${tag}_${w}_${v}.err
    python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/${tag}_${w}_${v}.json").read().strip().splitlines()[-1])
    r=j["roofline"]; pl=(r.get("hbm",r))["pipeline"]
    print("$w $v value %.1f Msamples/s  Mrays/s %.0f  ms_trace %.1f ms_shade %.1f"%(j["value"], j["extra"]["Mrays_per_s"], pl["ms_trace"], pl["ms_shade"]), j["extra"]["bvh"]["trace_layout"], "scene ms %.1f"%j["extra"]["bvh"]["total_ms"])
except Exception as e: print("$w $v failed", e)
PY
  done
done
unset MCPT_BVH8
