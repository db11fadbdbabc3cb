out=gpurun_out/v29; mkdir -p $out
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelin or path_mode or cfg1" 2>&1 | tail -3) > $out/pytest_pipe.log; cat $out/pytest_pipe.log
WORKLOAD=cfg1_default_4032_waters timeout 600 python scripts/profile_case.py 4032 20000 1 > $out/plain.log 2>&1; tail -3 $out/plain.log
MRISIM_PIPE_OVERLAP=0 WORKLOAD=cfg1_default_4032_waters timeout 600 python scripts/profile_case.py 4032 20000 1 > $out/plain_off.log 2>&1; tail -2 $out/plain_off.log
timeout 300 python scripts/crossover.py 8000 16384 32768 49152 > $out/crossover3.json 2> $out/crossover3.err; grep -c walkers $out/crossover3.json
