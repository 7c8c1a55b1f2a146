MCPT_DEBUG_TIMING=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload veach-mis > gpurun_out/dbg.json 2>gpurun_out/dbg.err
grep -n "bounce\|render_device" gpurun_out/dbg.err | head -80
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/ex_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/ex_launches.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
for r in rows[1:40]:
    print(r[ki][:60], r[vi])
PY
