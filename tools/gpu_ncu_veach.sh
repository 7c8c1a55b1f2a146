# usage: bash tools/gpu_ncu_veach.sh <tag> -- one `ncu --set full` capture (with SASS source counters) of the shading / selection /
# trace kernels of the headline workload at 32 spp per step, after a plain run of the same command has exited 0
tag=${1:-n}
timeout 300 python bench.py --steps 1 --warmup 0 --spp-per-step 32 --no-cpu-baseline --no-extras > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.err; exit 1; }
MCPT_BVH8=0 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_lightprep|k_lightsel|k_lsel|k_extend|k_gather" --launch-skip 0 --launch-count 16 -o gpurun_out/${tag}_veach_head -f python bench.py --steps 1 --warmup 0 --spp-per-step 32 --no-cpu-baseline --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1
ls -la gpurun_out/${tag}_veach_head.ncu-rep
