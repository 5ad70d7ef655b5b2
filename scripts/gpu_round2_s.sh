#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu.log | cut -c1-200
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_ch3oh_n1.json 2> gpurun_out/r02_bench_ch3oh_n1.err; echo "bench ch3oh rc=$?"
python bench.py --config h2o --steps 5 --warmup 3 > gpurun_out/r02_bench_h2o_n1.json 2> gpurun_out/r02_bench_h2o_n1.err; echo "bench h2o rc=$?"
python bench.py --batch 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_batch4_n1.json 2> gpurun_out/r02_bench_ch3oh_batch4_n1.err; echo "bench batch4 rc=$?"
python bench.py --start chained --batch 8 --layers 40 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch8_L40.json 2> /dev/null; echo "bench chained8 rc=$?"
python bench.py --start chained --batch 1 --layers 40 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02_bench_ch3oh_chained_batch1_L40.json 2> /dev/null; echo "bench chained1 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 700 --csv --log-file gpurun_out/r02_ncu_launches_bench_ch3oh.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-grid 0 > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python - <<'PY'
import json
for f in ['ch3oh_n1','h2o_n1','ch3oh_batch4_n1','ch3oh_chained_batch8_L40','ch3oh_chained_batch1_L40']:
    d=json.load(open(f'gpurun_out/r02_bench_{f}.json'))
    print(f, round(d['value'],3), round(d['e2e']['value'],3), round(d['ms_per_step'],1), d['roofline']['kernel'], round(d['roofline']['frac'],3), (round(d['sweep']['value'],2), d['sweep']['answer_digest']) if d.get('sweep') else '', {k:round(v['ms_per_step'],1) for k,v in d['kernels'].items() if k.startswith('lu') or k=='assemble'})
PY
