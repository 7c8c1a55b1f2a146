# usage: bash tools/gpu_launches.sh <tag> [bench args] -- ncu launch list (gpu__time_duration) of one bench step + per-kernel summary
tag=${1:-l}
shift
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras "$@" > gpurun_out/${tag}_launches.log 2>&1
python tools/launch_summary.py gpurun_out/${tag}_launches.csv > gpurun_out/${tag}_launches.summary.txt 2>&1; cat gpurun_out/${tag}_launches.summary.txt
