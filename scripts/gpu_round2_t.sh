#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_slab.py tests/test_adapter.py -m gpu -x -q --tb=short -p no:cacheprovider -k "edge or h2o or odd" > gpurun_out/r02t_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r02t_pytest.log | cut -c1-300
python bench.py --config h2o150 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02t_bench_h2o150_reg.json 2> gpurun_out/r02t_reg.err; echo "bench reg rc=$?"
RT_B200_FUSED_SHARED_LU=1 python bench.py --config h2o150 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02t_bench_h2o150_shared.json 2> gpurun_out/r02t_sh.err; echo "bench shared rc=$?"
python - <<'PY'
import json
for f in ['reg','shared']:
    try:
        d=json.load(open(f'gpurun_out/r02t_bench_h2o150_{f}.json'))
        print(f, round(d['value'],3), round(d['e2e']['value'],3), round(d['ms_per_step'],2))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/r02t_reg.err
