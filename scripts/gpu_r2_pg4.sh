#!/bin/bash
# pipelined path after the launch-group / draw-chunk retune: parity, timings over the walker count, crossover against the fused path
cd "$(dirname "$0")/.."
out=gpurun_out/pg5; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "pipelin or path_mode or full_default" > $out/pytest.txt 2>&1; tail -2 $out/pytest.txt
python scripts/crossover.py 4032 8000 16384 32768 49152 65536 81920 > $out/crossover.json 2> $out/err.log
python - <<'PY'
import json
for r in json.load(open('gpurun_out/pg5/crossover.json')):
    if isinstance(r, dict) and 'walkers' in r: print(r['walkers'], r['path_mode'], round(r['ms_per_sprint'], 2))
PY
