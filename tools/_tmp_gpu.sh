cp monte_carlo_path_tracing_b200/libmcpt.so /tmp/libmcpt_base.so
for v in base ct2.0f ct4.0f; do
  if [ $v = base ]; then cp /tmp/libmcpt_base.so monte_carlo_path_tracing_b200/libmcpt.so; else cp build_variants/libmcpt_$v.so monte_carlo_path_tracing_b200/libmcpt.so; fi
  for w in soup-1m interior veach-mis; do
  timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --workload $w 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); p=j['roofline']['pipeline']; print('$v $w', round(j['value'],1), 'trace', round(p['ms_trace']/2,2), 'sah', round(j['extra']['bvh']['sah'],1), 'nodes', j['extra']['bvh']['nodes'])"
  done
done
cp /tmp/libmcpt_base.so monte_carlo_path_tracing_b200/libmcpt.so
