out=gpurun_out/v34; mkdir -p $out
(timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > $out/pytest_gpu.txt; cat $out/pytest_gpu.txt
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; tail -1 $out/smoke.log
