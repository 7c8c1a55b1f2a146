timeout 600 python -m pytest tests -m gpu -x -q -k "gpu_bvh or checkpoint or cpp_driver" -s 2>&1 | tail -12
for b in host gpu; do
timeout 300 python bench.py --steps 3 --warmup 3 --workload soup-1m --no-cpu-baseline --bvh $b > gpurun_out/r1u_bench_soup_$b.json 2> gpurun_out/r1u_bench_soup_$b.err; tail -c 300 gpurun_out/r1u_bench_soup_$b.err
python - <<PY
import json
j=json.load(open("gpurun_out/r1u_bench_soup_$b.json")); print("$b", j["value"], j["e2e"]["value"], j["extra"]["bvh"])
PY
done
