python bench.py --steps 3 --warmup 3 --workload cornell-box --spp-per-step 64 --batch-spp 16 --no-cpu-baseline > gpurun_out/r1b_cornell_b16.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1b_launches_veach.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1b_launches_cornell.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload cornell-box --spp-per-step 64 > /dev/null 2>&1
