mkdir -p gpurun_out/v6c
(python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/v6c/pytest.log; tail -3 gpurun_out/v6c/pytest.log
python scripts/profile_case.py 1048576 1000 1 > gpurun_out/v6c/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:walker_sprint -s 81 -c 1 -o gpurun_out/v6c/prof python scripts/profile_case.py 1048576 1000 1 > gpurun_out/v6c/ncu.log 2>&1
cat gpurun_out/v6c/plain.log
