#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/time_solve.py ch3oh 8 2 2>&1 | tail -1 | cut -c1-200
timeout 200 python scripts/time_solve.py ch3oh 200 2 2>&1 | tail -1 | cut -c1-200
timeout 200 python scripts/time_solve.py h2o 200 2 2>&1 | tail -1 | cut -c1-200
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_gpu.log | cut -c1-200
