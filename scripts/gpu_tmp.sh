#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_slab.py -m gpu -x -q --tb=short -p no:cacheprovider -k "not ch3oh" > gpurun_out/tmp_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/tmp_pytest.log | cut -c1-200
for cfg in oh oh_overlap h2o150; do
timeout 300 python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', round(d['value'],3), round(d['ms_per_step'],2))"
done
