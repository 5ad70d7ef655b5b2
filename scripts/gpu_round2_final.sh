#!/bin/bash
# Final single-GPU evidence of round 2 (after the PDL / stream-priority changes): GPU suite, bench lines, launch list
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu.log | cut -c1-300
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_ch3oh_n1.json 2> gpurun_out/r02_bench_ch3oh_n1.err; echo "bench ch3oh rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ch3oh_reference_arm.json 2> gpurun_out/r02_bench_ch3oh_reference_arm.err; echo "ref arm rc=$?"
for cfg in h2o h2o150 oh oh_overlap; do
  python bench.py --config $cfg --steps 5 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/r02_bench_${cfg}_n1.err; echo "bench $cfg rc=$?"
done
python bench.py --batch 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_batch4_n1.json 2> gpurun_out/r02_bench_ch3oh_batch4_n1.err; echo "bench batch4 rc=$?"
python bench.py --start chained --batch 8 --layers 40 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch8_L40.json 2> gpurun_out/r02_bench_ch3oh_chained_batch8_L40.err; echo "bench chained rc=$?"
python bench.py --start chained --batch 1 --layers 40 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch1_L40.json 2> gpurun_out/r02_bench_ch3oh_chained_batch1_L40.err; echo "bench chained1 rc=$?"
python bench.py --cpu-full-slab --steps 3 --warmup 3 --sweep-grid 0 > gpurun_out/r02_bench_ch3oh_n1_cpu_full_slab.json 2> gpurun_out/r02_bench_ch3oh_n1_cpu_full_slab.err; echo "bench cpu full slab rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 700 --csv --log-file gpurun_out/r02_ncu_launches_bench_ch3oh.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-grid 0 > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
RT_B200_LU_LOOKAHEAD=0 RT_B200_LU_GROUPS=1 ncu --set full --import-source on --clock-control none -k regex:"k_panel|k_trsm|k_backsolve|k_assemble" -s 60 -c 8 -o gpurun_out/r02_lu_others -f python scripts/profile_step.py ch3oh 200 1 > gpurun_out/r02_ncu_lu_others.log 2>&1; echo "ncu others rc=$?"
ncu --set full --import-source on --clock-control none -k regex:k_fused -c 1 -o gpurun_out/r02_fused_h2o150 -f python scripts/profile_solve.py h2o150 200 > gpurun_out/r02_ncu_fused.log 2>&1; echo "ncu fused rc=$?"
for f in gpurun_out/r02_bench_*_n1*.json; do echo $f; cut -c1-160 $f; done
