timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
timeout 600 python tools/strategy_sweep.py --out gpurun_out/r1u_strategy_sweep.json > gpurun_out/r1u_strategy_sweep.log 2>&1; tail -12 gpurun_out/r1u_strategy_sweep.log
