#!/bin/bash
# correctness first (small parity cases wrapped in timeouts), then timings of the paired-prefix kernel
cd "$(dirname "$0")/.."
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tri or work_plan or reference_cases" 2>&1 | tail -5
timeout 120 python scripts/run_one.py cpx64 8 5
timeout 120 python scripts/run_one.py c4 8 5
timeout 120 python scripts/run_one.py cpx 8 5
timeout 200 python scripts/tri_sweep.py 40 48 56 64 72 80 2>&1 | grep "kernel tri"
