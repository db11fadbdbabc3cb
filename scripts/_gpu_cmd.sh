out=gpurun_out/v36; mkdir -p $out
timeout 50 python bench.py --workload cfg2_8000_waters_intracellular_kick --no-cpu-baseline > $out/bench_cfg2.json 2> $out/bench_cfg2.err; cut -c1-200 $out/bench_cfg2.json
