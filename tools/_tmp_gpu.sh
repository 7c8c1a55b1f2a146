timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "soup_closest" 2>&1 | grep -E "^E|assert" | head -20
