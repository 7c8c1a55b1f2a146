# usage: bash tools/gpu_ab_edge.sh <tag> [variant suffix] -- GPU tests on the shipped library, then the headline with the shipped and one variant library, twice
tag=${1:-e}
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${tag}_pytest_gpu.log
bash tools/ab_libs.sh veach-mis - ${2:-_noedge} - ${2:-_noedge}
