timeout 900 python -m pytest tests -m gpu -x -q -k "tiny or without_lights or veach_full" 2>&1 | tail -15
