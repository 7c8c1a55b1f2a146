# usage: bash tools/gpu_check_big.sh <tag>  -- the two large-scene workloads (configs[2], configs[3])
tag=${1:-run}
timeout 300 python bench.py --steps 3 --warmup 3 --workload soup-1m --no-cpu-baseline > gpurun_out/${tag}_bench_soup.json 2> gpurun_out/${tag}_bench_soup.err; tail -c 400 gpurun_out/${tag}_bench_soup.err
timeout 300 python bench.py --steps 3 --warmup 3 --workload interior --no-cpu-baseline > gpurun_out/${tag}_bench_interior.json 2> gpurun_out/${tag}_bench_interior.err; tail -c 400 gpurun_out/${tag}_bench_interior.err
python - <<PY
import json
for w in ("soup","interior"):
    try:
        j=json.loads(open("gpurun_out/${tag}_bench_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "value %.1f Msamples/s  ms/step %.2f  e2e %.1f  Msb/s %.0f  Mrays/s %.0f launches %d"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j["extra"]["Msample_bounces_per_s"],j["extra"]["Mrays_per_s"],j["gpu_launches"]))
    except Exception as e: print(w,"failed",e)
PY
