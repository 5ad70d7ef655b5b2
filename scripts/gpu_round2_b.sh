#!/bin/bash
# GPU call B of round 2: no-alloc entry point alone in a fresh process, then the whole GPU suite (the no-alloc test runs last)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/r02b_pytest_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?"; tail -3 gpurun_out/r02b_pytest_noalloc_alone.log
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/r02b_pytest_gpu.log 2>&1
echo "pytest all rc=$?"; tail -15 gpurun_out/r02b_pytest_gpu.log | cut -c1-300
compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02b_sanitizer_noalloc.log 2>&1
echo "sanitizer rc=$?"; tail -4 gpurun_out/r02b_sanitizer_noalloc.log
