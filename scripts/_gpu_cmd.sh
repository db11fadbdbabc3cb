mkdir -p gpurun_out/v7
(python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/v7/pytest.log; tail -3 gpurun_out/v7/pytest.log
RESORTS="250 125" bash scripts/tune_variants.sh 2>&1 | tee gpurun_out/v7/tune.log
