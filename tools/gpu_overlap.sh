# usage: bash tools/gpu_overlap.sh -- the two-stream overlap probe with full and with halved persistent grids
for d in 1 2; do echo "== MCPT_GRID_DIV=$d"; MCPT_GRID_DIV=$d timeout 600 python tools/diag/overlap_probe.py 0 4 8 12 2>&1 | tail -8; done
