#!/bin/bash
python scripts/profile_solve.py h2o150 200 2>&1 | tail -1
python scripts/profile_solve.py h2o150 200 2>&1 | tail -1
python -m pytest tests/test_gpu_full_slab.py -m gpu -q -k "h2o150" -p no:cacheprovider 2>&1 | tail -2
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "live_solve or h2o" -p no:cacheprovider 2>&1 | tail -2
