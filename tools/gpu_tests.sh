# usage: bash tools/gpu_tests.sh <tag> [pytest args]  -- the GPU test suite + smoke, logs under gpurun_out/<tag>_*
tag=${1:-run}; shift
timeout 1500 python -m pytest tests -m gpu -q --durations=15 "$@" > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${tag}_smoke.log
