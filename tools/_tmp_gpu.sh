# scratch script for one-off gpurun experiments (overwritten freely; see tools/gpu_round.sh for the round-end checks)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
