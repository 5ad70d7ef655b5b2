#!/bin/bash
# final measurements of the small-molecule path after the register-resident LU was extended to every N <= 150
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for cfg in oh oh_overlap h2o150; do
timeout 400 python bench.py --config $cfg --steps 5 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/r02_bench_${cfg}_n1.err; echo "bench $cfg rc=$?"
done
timeout 300 python bench.py --config oh --batch 3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_batch3_n1.json 2> /dev/null; echo "bench oh batch3 rc=$?"
timeout 300 python bench.py --config oh_overlap --batch 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_oh_overlap_batch2_n1.json 2> /dev/null; echo "bench oh_overlap batch2 rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_fused -c 1 -o gpurun_out/r02_fused_h2o150 -f python scripts/profile_solve.py h2o150 200 > gpurun_out/r02_ncu_fused.log 2>&1; echo "ncu fused h2o150 rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_fused -c 1 -o gpurun_out/r02_fused_oh -f python scripts/profile_solve.py oh 100 > gpurun_out/r02_ncu_fused_oh.log 2>&1; echo "ncu fused oh rc=$?"
python - <<'PY'
import json
for f in ['oh_n1','oh_overlap_n1','h2o150_n1','oh_batch3_n1','oh_overlap_batch2_n1']:
    try:
        d=json.load(open(f'gpurun_out/r02_bench_{f}.json'))
        print(f, round(d['value'],3), round(d['e2e']['value'],3), round(d['ms_per_step'],2), d.get('cpu_baseline',{}).get('value'), d['roofline'].get('kernel'), round(d['roofline'].get('frac') or 0,3))
    except Exception as e: print(f, 'ERR', e)
PY
