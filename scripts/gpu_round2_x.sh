#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_slab.py -m gpu -x -q --tb=short -p no:cacheprovider -k "not ch3oh" > gpurun_out/r02x_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r02x_pytest.log | cut -c1-300
for cfg in oh oh_overlap h2o150; do
timeout 300 python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02x_bench_${cfg}.json 2> gpurun_out/r02x_${cfg}.err; echo "bench $cfg rc=$?"
done
python - <<'PY'
import json
for f in ['oh','oh_overlap','h2o150']:
    try:
        d=json.load(open(f'gpurun_out/r02x_bench_{f}.json'))
        print(f, round(d['value'],3), round(d['e2e']['value'],3), round(d['ms_per_step'],2))
    except Exception as e: print(f, 'ERR', e)
PY
