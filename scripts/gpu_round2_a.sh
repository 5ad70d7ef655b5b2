#!/bin/bash
# GPU call A of round 2: the whole GPU test suite (incl. the new all-layer parity tests), bench lines, panel ncu capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/r02a_pytest_gpu.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/r02a_pytest_gpu.log
tail -30 gpurun_out/r02a_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench_ch3oh_n1.json 2> gpurun_out/r02a_bench_ch3oh_n1.err; echo "bench ch3oh rc=$?"
python bench.py --config h2o --steps 3 --warmup 3 > gpurun_out/r02a_bench_h2o_n1.json 2> gpurun_out/r02a_bench_h2o_n1.err; echo "bench h2o rc=$?"
python bench.py --config h2o150 --steps 3 --warmup 3 > gpurun_out/r02a_bench_h2o150_n1.json 2> gpurun_out/r02a_bench_h2o150_n1.err; echo "bench h2o150 rc=$?"
ncu --set full --import-source on --clock-control none -k regex:k_panel -s 0 -c 1 -o gpurun_out/r02a_panel -f python scripts/profile_step.py ch3oh 200 1 > gpurun_out/r02a_ncu_panel.log 2>&1; echo "ncu panel rc=$?"
ncu --set full --import-source on --clock-control none -k regex:k_trsm -s 1 -c 1 -o gpurun_out/r02a_trsm -f python scripts/profile_step.py ch3oh 200 1 > gpurun_out/r02a_ncu_trsm.log 2>&1; echo "ncu trsm rc=$?"
cut -c1-400 gpurun_out/r02a_bench_ch3oh_n1.json
