#!/bin/bash
# quick GPU check: parity tests for the forward paths + decode/train timings
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/gpu_perf.py 64 16384 2>&1 | head -4
