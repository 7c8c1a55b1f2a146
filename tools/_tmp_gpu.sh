timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload interior > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1w_launches_interior.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload interior > /dev/null 2>&1
ls -la gpurun_out/r1w_launches_interior.csv
