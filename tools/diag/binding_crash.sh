python - <<PY
import sys; sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import util
sc = util.scene_text("cornell-box"); print(sc.write("/tmp/bc"))
PY
cp monte_carlo_path_tracing_b200/libmcpt.so /tmp/bc/libmcpt_orig.so
cp monte_carlo_path_tracing_b200/libmcpt_dbg.so monte_carlo_path_tracing_b200/libmcpt.so
for s in nee_uniform mis; do
  oracle/_ref/ref_gpu_main /tmp/bc/cornell-box.obj /tmp/bc/cornell-box.xml /tmp/bc/o_$s.pfm --spp 8 --strategy $s --seed 3 --max-depth 6 > /tmp/bc/log_$s.txt 2>&1; echo "$s rc=$?"; strings /tmp/bc/log_$s.txt | tail -25
done
cp /tmp/bc/libmcpt_orig.so monte_carlo_path_tracing_b200/libmcpt.so
