# usage: WORKLOAD=veach-mis bash tools/gpu_scale.sh <tag> <N...>  -- bench.py at each N on one box (torchrun for N > 1)
wl=${WORKLOAD:-veach-mis}
tag=$1; shift
for n in "$@"; do
  if [ "$n" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --steps 3 --warmup 3 --workload $wl --no-cpu-baseline > gpurun_out/${tag}_scale_n1.json 2> gpurun_out/${tag}_scale_n1.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 --workload $wl > gpurun_out/${tag}_scale_n$n.json 2> gpurun_out/${tag}_scale_n$n.err
  fi
  tail -c 300 gpurun_out/${tag}_scale_n$n.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/${tag}_scale_n$n.json").read().strip().splitlines()[-1])
    print("N=$n value %.1f Msamples/s ms/step %.2f e2e %.1f"%(j["value"],j["ms_per_step"],j["e2e"]["value"]))
except Exception as e: print("N=$n failed",e)
PY
done
