python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in "veach-mis 16" "veach-mis 64" "interior 8" "interior 32" "cornell-box 64" "cornell-box 256" "soup-1m 16" "soup-1m 64"; do
  set -- $cfg
  python bench.py --steps 3 --warmup 3 --workload $1 --spp-per-step $2 --no-cpu-baseline > gpurun_out/tmp.json 2> gpurun_out/tmp.err || tail -3 gpurun_out/tmp.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/tmp.json").read().strip().splitlines()[-1])
print("$1 spp/step $2: value %.1f Msamples/s  ms/step %.2f  e2e %.1f  Msb/s %.0f launches %d"%(j["value"],j["ms_per_step"],j["e2e"]["value"],j["extra"]["Msample_bounces_per_s"],j["gpu_launches"]))
PY
done
