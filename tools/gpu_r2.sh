# usage: bash tools/gpu_r2.sh <tag> [nopytest] -- round-2 GPU check: tests, smoke, default bench (+extras, cpu baseline), reference arm,
# the other workloads, then ncu: launch list + one --set full capture (with SASS source counters) of the shading / trace kernels.
tag=${1:-run}
if [ "$2" != "nopytest" ]; then
timeout 1500 python -m pytest tests -m gpu -q --durations=10 > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log
fi
timeout 900 python bench.py > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err; echo "bench rc=$?"; tail -c 400 gpurun_out/${tag}_bench_default.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
for w in cornell-box soup-1m interior; do
  timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
done
python - <<PY
import json
for w in ("default","reference","cornell-box","soup-1m","interior"):
    try:
        j=json.loads(open("gpurun_out/${tag}_bench_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "value %.2f  ms/step %.2f  e2e %.2f  launches %d"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j.get("gpu_launches",0)))
        if w=="default":
            print(" roofline", {k:j["roofline"][k] for k in ("bound","achieved","peak","frac","unit")}, "fp32", j["fp32"]["peak_source"])
            for k,v in j["extra"].items():
                if isinstance(v,dict) and k!="bvh": print("  ",k,v)
            print("  cpu", j.get("cpu_baseline"))
    except Exception as e: print(w,"failed",e)
PY
if [ "$3" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_veach_mis.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > /dev/null 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_lightprep|k_lightsel|k_extend|k_gather" --launch-skip 0 --launch-count 8 -o gpurun_out/${tag}_veach_head -f python bench.py --steps 1 --warmup 0 --spp-per-step 32 --no-cpu-baseline --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1
fi
ls gpurun_out | grep ${tag}_ | wc -l
