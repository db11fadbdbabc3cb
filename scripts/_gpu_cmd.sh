mkdir -p gpurun_out/v9
(python -m pytest tests -m "gpu and not slow" -x -q 2>&1 | tail -25) > gpurun_out/v9/pytest.log; tail -25 gpurun_out/v9/pytest.log
