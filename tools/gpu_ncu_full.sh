set -x
ncu --set full --clock-control none --import-source on -k regex:k_bounce --launch-skip 0 --launch-count 2 -o gpurun_out/r1_veach_mis_head -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_ncu_head_veach.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bounce --launch-skip 0 --launch-count 3 -o gpurun_out/r1_cornell_nee_head -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload cornell-box --spp-per-step 64 > gpurun_out/r1_ncu_head_cornell.log 2>&1
ls -la gpurun_out
