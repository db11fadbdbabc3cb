out=gpurun_out/v28; mkdir -p $out
MRISIM_LONG_XCHECK=1 MRISIM_XCHECK_OUT=$out timeout 900 python -m pytest tests/test_ref_gpu_crosscheck.py -m gpu -x -q -rP -k "default" 2>&1 | grep -v "^$" | tail -25 | cut -c1-400 > $out/xcheck_default.txt; cat $out/xcheck_default.txt
timeout 300 python scripts/crossover.py 40960 49152 57344 > $out/crossover2.json 2> $out/crossover2.err; grep -c walkers $out/crossover2.json
