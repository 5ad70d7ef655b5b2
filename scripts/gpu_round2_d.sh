#!/bin/bash
# GPU call D: whole GPU suite with the final fixtures, single-GPU bench lines of every config -> profiles/, ncu launch list +
# full captures of the trailing update (both staging variants) with the DMMA pipe counters
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/r02d_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -12 gpurun_out/r02d_pytest_gpu.log | cut -c1-300
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_ch3oh_n1.json 2> gpurun_out/r02_bench_ch3oh_n1.err; echo "bench ch3oh rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ch3oh_reference_arm.json 2> gpurun_out/r02_bench_ch3oh_reference_arm.err; echo "ref arm rc=$?"
for cfg in h2o h2o150 oh oh_overlap; do
  python bench.py --config $cfg --steps 3 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/r02_bench_${cfg}_n1.err; echo "bench $cfg rc=$?"
done
python bench.py --batch 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_batch4_n1.json 2> gpurun_out/r02_bench_ch3oh_batch4_n1.err; echo "bench batch4 rc=$?"
python bench.py --start chained --batch 8 --layers 40 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch8_L40.json 2> gpurun_out/r02_bench_ch3oh_chained_batch8_L40.err; echo "bench chained rc=$?"
python bench.py --start chained --batch 1 --layers 40 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch1_L40.json 2> gpurun_out/r02_bench_ch3oh_chained_batch1_L40.err; echo "bench chained1 rc=$?"
python bench.py --config oh --batch 3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_batch3_n1.json 2> gpurun_out/r02_bench_oh_batch3_n1.err; echo "bench oh batch3 rc=$?"
python bench.py --config oh_overlap --batch 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_overlap_batch2_n1.json 2> gpurun_out/r02_bench_oh_overlap_batch2_n1.err; echo "bench oh_overlap batch2 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 600 --csv --log-file gpurun_out/r02_ncu_launches_bench_ch3oh.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-grid 0 > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for mode in 3 4; do
RT_B200_GEMM_MODE=$mode RT_B200_LU_LOOKAHEAD=0 RT_B200_LU_GROUPS=1 ncu --set full --import-source on --clock-control none -k regex:k_gemm -s 23 -c 3 -o gpurun_out/r02_gemm_mode$mode -f python scripts/profile_step.py ch3oh 200 1 > gpurun_out/r02_ncu_gemm$mode.log 2>&1; echo "ncu gemm mode $mode rc=$?"
done
for f in gpurun_out/r02_bench_*_n1.json gpurun_out/r02_bench_ch3oh_chained*.json; do echo $f; cut -c1-220 $f; done
