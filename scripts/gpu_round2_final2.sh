#!/bin/bash
# final evidence of round 2 with the final tree: GPU suite, smoke, headline and small-molecule bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_ch3oh_n1.json 2> gpurun_out/r02_bench_ch3oh_n1.err; echo "bench ch3oh rc=$?"
timeout 400 python bench.py --config h2o --steps 5 --warmup 3 > gpurun_out/r02_bench_h2o_n1.json 2> gpurun_out/r02_bench_h2o_n1.err; echo "bench h2o rc=$?"
for cfg in oh oh_overlap h2o150; do
timeout 400 python bench.py --config $cfg --steps 5 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/r02_bench_${cfg}_n1.err; echo "bench $cfg rc=$?"
done
timeout 300 python bench.py --config oh --batch 3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_batch3_n1.json 2> /dev/null; echo "bench oh batch3 rc=$?"
timeout 300 python bench.py --config oh_overlap --batch 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_overlap_batch2_n1.json 2> /dev/null; echo "bench oh_overlap batch2 rc=$?"
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02_bench_ch3oh_reference_arm.json 2> /dev/null; echo "reference arm rc=$?"
python - <<'PY'
import json
for f in ['ch3oh_n1','h2o_n1','oh_n1','oh_overlap_n1','h2o150_n1','oh_batch3_n1','oh_overlap_batch2_n1','ch3oh_reference_arm']:
    try:
        d=json.load(open(f'gpurun_out/r02_bench_{f}.json'))
        print(f, round(d['value'],4), round(d['e2e']['value'],4), round(d['ms_per_step'],2), (d.get('cpu_baseline') or {}).get('value'), (d.get('roofline') or {}).get('kernel'), round((d.get('roofline') or {}).get('frac') or 0,3), (round(d['sweep']['value'],2), d['sweep']['answer_digest']) if d.get('sweep') else '', d.get('clocks'))
    except Exception as e: print(f, 'ERR', e)
PY
