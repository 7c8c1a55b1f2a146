# usage: bash tools/gpu_ab_all.sh <tag> <suffix> [<suffix> ...] -- GPU tests on the shipped library, then every workload with the shipped ("-") and the variant libraries
tag=$1; shift
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
for w in veach-mis interior soup-1m cornell-box; do bash tools/ab_libs.sh $w - "$@"; done
