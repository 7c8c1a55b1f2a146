# usage: bash tools/gpu_check.sh <tag>   -- GPU parity tests + the two bench workloads, outputs under gpurun_out/<tag>_*
tag=${1:-run}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${tag}_pytest_gpu.log
tail -3 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_veach.json 2> gpurun_out/${tag}_bench_veach.err
timeout 300 python bench.py --steps 3 --warmup 3 --workload cornell-box --no-cpu-baseline > gpurun_out/${tag}_bench_cornell.json 2> gpurun_out/${tag}_bench_cornell.err
tail -c 600 gpurun_out/${tag}_bench_veach.err gpurun_out/${tag}_bench_cornell.err
python - <<PY
import json
for w in ("veach","cornell"):
    try:
        j=json.load(open("gpurun_out/${tag}_bench_%s.json"%w))
        print(w, "value %.1f Msamples/s  ms/step %.2f  e2e %.1f  Msb/s %.0f  Mrays/s %.0f launches %d"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j["extra"]["Msample_bounces_per_s"],j["extra"]["Mrays_per_s"],j["gpu_launches"]))
    except Exception as e: print(w,"failed",e)
PY
