# scratch script for one-off gpurun experiments (overwritten freely; see tools/gpu_round.sh for the round-end checks)
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(round(j['value'],1), j['fp32']['frac'], j['roofline']['frac'], j['gpu_launches'], j['clocks'])"
