# scratch script for one-off gpurun experiments (overwritten freely; see tools/gpu_round.sh for the round-end checks)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for w in veach-mis interior; do
timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --workload $w 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); p=j['roofline']['pipeline']; print('$w', round(j['value'],1), 'shade', round(p['ms_shade']/2,2), 'trace', round(p['ms_trace']/2,2))"
done
MCPT_DEBUG_TIMING=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload veach-mis --light-tree 2>gpurun_out/rb.err | python -c "
import json,sys; j=json.loads(sys.stdin.read()); print('veach tree', round(j['value'],1))"; grep "traced again" gpurun_out/rb.err | head -2
