# usage: bash tools/gpu_round.sh <tag>  -- everything a round-end check needs, outputs under gpurun_out/<tag>_*
# GPU tests, smoke, the four bench workloads, the default bench line (with cpu_baseline), the reference arm,
# then (only after the plain runs exited 0) the ncu launch list and one --set full capture of the dominant kernels.
tag=${1:-run}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${tag}_pytest_gpu.log; tail -2 gpurun_out/${tag}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
for w in veach-mis cornell-box soup-1m interior; do
  timeout 300 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
done
timeout 300 python bench.py --steps 3 --warmup 3 --workload veach-mis --light-tree --no-cpu-baseline > gpurun_out/${tag}_bench_veach-mis-lighttree.json 2> gpurun_out/${tag}_bench_veach-mis-lighttree.err
timeout 400 python bench.py > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
python - <<PY
import json
for w in ("veach-mis","cornell-box","soup-1m","interior","veach-mis-lighttree","default","reference"):
    try:
        j=json.loads(open("gpurun_out/${tag}_bench_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "value %.2f  ms/step %.2f  e2e %.2f  launches %d"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j.get("gpu_launches",0)), j.get("extra",{}))
    except Exception as e: print(w,"failed",e)
PY
if [ "$2" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_veach_mis.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_lightprep|k_lightsel|k_extend|k_gather" --launch-skip 0 --launch-count 8 -o gpurun_out/${tag}_veach_head -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
fi
if [ "$2" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_veach_mis_lighttree.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --light-tree > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_lighttree" --launch-skip 0 --launch-count 2 -o gpurun_out/${tag}_veach_lighttree -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --light-tree > /dev/null 2>&1
fi
ls gpurun_out | grep ${tag}_ | wc -l
