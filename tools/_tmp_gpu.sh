cp monte_carlo_path_tracing_b200/libmcpt.so /tmp/libmcpt_base.so
for v in base u1 u3 u4 c512; do
  if [ $v = base ]; then cp /tmp/libmcpt_base.so monte_carlo_path_tracing_b200/libmcpt.so; else cp build_variants/libmcpt_$v.so monte_carlo_path_tracing_b200/libmcpt.so; fi
  timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --workload veach-mis 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); p=j['roofline']['pipeline']; print('$v veach', round(j['value'],1), 'shade', round(p['ms_shade']/3,2), 'trace', round(p['ms_trace']/3,2))"
done
cp build_variants/libmcpt_c512.so monte_carlo_path_tracing_b200/libmcpt.so
for w in cornell-box interior; do timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --workload $w 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); print('c512 $w', round(j['value'],1))"; done
cp /tmp/libmcpt_base.so monte_carlo_path_tracing_b200/libmcpt.so
