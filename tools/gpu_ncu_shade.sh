# usage: bash tools/gpu_ncu_shade.sh <tag> <workload> -- one `ncu --set full` capture (SASS source counters) of the shading-side kernels of a workload
tag=${1:-t}; w=${2:-interior}
timeout 300 python bench.py --steps 1 --warmup 0 --spp-per-step 32 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.err; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_lightsel|k_gather|k_runsum|k_lightprep" --launch-skip 0 --launch-count 10 -o gpurun_out/${tag}_${w}_shade -f python bench.py --steps 1 --warmup 0 --spp-per-step 32 --workload $w --no-cpu-baseline --no-extras > gpurun_out/${tag}_ncu_shade.log 2>&1
ls -la gpurun_out/${tag}_${w}_shade.ncu-rep
