out=gpurun_out/v32; mkdir -p $out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > $out/pytest_gpu.txt; cat $out/pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; tail -1 $out/smoke.log
python bench.py --workload cfg1_default_4032_waters --no-cpu-baseline > $out/bench_cfg1.json 2> $out/bench_cfg1.err; cut -c1-200 $out/bench_cfg1.json
python bench.py --workload cfg2_8000_waters_intracellular_kick --no-cpu-baseline > $out/bench_cfg2.json 2> $out/bench_cfg2.err; cut -c1-200 $out/bench_cfg2.json
