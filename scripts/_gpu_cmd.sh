out=gpurun_out/v24; mkdir -p $out
export WORKLOAD=cfg1_default_4032_waters
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelin" 2>&1 | tail -3) > $out/pytest_pipe.log; cat $out/pytest_pipe.log
timeout 600 python scripts/profile_case.py 4032 20000 1 > $out/plain.log 2>&1; tail -3 $out/plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"walker_path" -s 2 -c 1 -o /tmp/p2 \
    python scripts/profile_case.py 4032 20000 1 > $out/ncu.log 2>&1
SORT=samples python scripts/ncu_lines.py /tmp/p2.ncu-rep 12096000 40 > $out/p2_lines_by_samples.txt 2>&1
ncu -i /tmp/p2.ncu-rep --page details > $out/p2_details.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"walker_|step_field|chunk_states" -s 40 -c 40 --csv --log-file $out/launches_pipe.csv python scripts/profile_case.py 4032 20000 1 > /dev/null 2>&1
