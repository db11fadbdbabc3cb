out=gpurun_out/v30; mkdir -p $out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > $out/pytest_gpu.txt; cat $out/pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; tail -1 $out/smoke.log
python bench.py --workload cfg3_100k_waters_low_permeability --no-cpu-baseline > $out/bench_cfg3.json 2> $out/bench_cfg3.err; cut -c1-200 $out/bench_cfg3.json
