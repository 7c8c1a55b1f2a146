set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/r1_smoke.log 2>&1
python bench.py --steps 3 --warmup 3 > gpurun_out/r1_bench_veach.json 2> gpurun_out/r1_bench_veach.err
python bench.py --steps 3 --warmup 3 --workload cornell-box --spp-per-step 64 --no-cpu-baseline > gpurun_out/r1_bench_cornell.json 2> gpurun_out/r1_bench_cornell.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_veach.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_ncu_list.log 2>&1
tail -3 gpurun_out/*.log gpurun_out/*.json
