#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02f_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/r02f_pytest_gpu.log | cut -c1-300
for cfg in h2o h2o150; do
  python bench.py --config $cfg --steps 3 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/r02_bench_${cfg}_n1.err; echo "bench $cfg rc=$?"
  cut -c1-200 gpurun_out/r02_bench_${cfg}_n1.json
done
python bench.py --steps 3 --warmup 3 --sweep-batch 4 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_n1_sweep_batch4.json 2> gpurun_out/r02_bench_ch3oh_n1_sweep_batch4.err; echo "bench sweep batch4 rc=$?"
python bench.py --steps 3 --warmup 3 --sweep-batch 8 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_n1_sweep_batch8.json 2> gpurun_out/r02_bench_ch3oh_n1_sweep_batch8.err; echo "bench sweep batch8 rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
