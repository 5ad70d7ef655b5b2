#!/bin/bash
# GPU call C: trailing-update variants (C staged by cp.async), fused N=150 check, ncu --set full of the gemm with every metric
mkdir -p gpurun_out
for mode in 3 4; do
  RT_B200_GEMM_MODE=$mode python scripts/time_step.py ch3oh 200 5 2>&1 | tail -2
  RT_B200_GEMM_MODE=$mode python scripts/time_step.py ch3oh 60 5 2>&1 | tail -2
done
for mode in 3 4; do
  RT_B200_GEMM_MODE=$mode python scripts/time_solve.py ch3oh 200 3 2>&1 | tail -2
done
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "live_solve or golden_converged" -p no:cacheprovider 2>&1 | tail -4
python -m pytest tests/test_gpu_full_slab.py -m gpu -q -k "h2o or oh" -p no:cacheprovider 2>&1 | tail -6 | cut -c1-300
python bench.py --config h2o150 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_bench_h2o150_n1.json 2> gpurun_out/r02c_bench_h2o150_n1.err; echo "bench h2o150 rc=$?"
cut -c1-300 gpurun_out/r02c_bench_h2o150_n1.json
for mode in 3 4; do
RT_B200_GEMM_MODE=$mode ncu --set full --import-source on --clock-control none -k regex:k_gemm -s 1 -c 1 -o gpurun_out/r02c_gemm_mode$mode -f python scripts/profile_step.py ch3oh 200 1 > gpurun_out/r02c_ncu_gemm$mode.log 2>&1; echo "ncu gemm mode $mode rc=$?"
done
