out=gpurun_out/v33; mkdir -p $out
export WORKLOAD=cfg1_default_4032_waters
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelin or path_mode or cfg1" 2>&1 | tail -3) > $out/pytest_pipe.log; cat $out/pytest_pipe.log
timeout 600 python scripts/profile_case.py 4032 20000 1 > $out/plain.log 2>&1; tail -3 $out/plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"walker_|step_field|chunk_s" -s 48 -c 48 --csv --log-file $out/launches_pipe.csv python scripts/profile_case.py 4032 20000 1 > /dev/null 2>&1
