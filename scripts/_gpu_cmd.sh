mkdir -p gpurun_out/oct4
(timeout 900 python -m pytest tests/test_octree_gpu.py -m "gpu" -x -q 2>&1 | tail -15) > gpurun_out/oct4/pytest.log; tail -4 gpurun_out/oct4/pytest.log
timeout 600 python scripts/sweep_config5.py --waters 1048576 --sprints 1 --mmoments 3.5e-17,7e-17 --stdev-factors 1,2 > gpurun_out/oct4/sweep_1gpu_small.jsonl 2> gpurun_out/oct4/sweep.err; cat gpurun_out/oct4/sweep_1gpu_small.jsonl; tail -3 gpurun_out/oct4/sweep.err
for wl in cfg1_default_4032_waters cfg2_8000_waters_intracellular_kick cfg3_lowperm_100k_waters; do for pm in auto fused; do
  timeout 600 python bench.py --workload $wl --path-mode $pm --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/oct4/bench_${wl}_${pm}.json 2>> gpurun_out/oct4/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/oct4/bench_${wl}_${pm}.json')); print('$wl','$pm',d['value'],d['e2e']['value'],d['ms_per_step'],d['config'].get('pipeline_sprints'),d['gpu_launches'])"
done; done
tail -3 gpurun_out/oct4/bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:leaf_field_kernel -s 8 -c 1 -o /tmp/leaf_prof python scripts/octree_bench.py --no-host dense2000 > gpurun_out/oct4/ncu_leaf.log 2>&1; tail -2 gpurun_out/oct4/ncu_leaf.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:criterion_kernel -s 8 -c 1 -o /tmp/crit_prof python scripts/octree_bench.py --no-host dense2000 > gpurun_out/oct4/ncu_crit.log 2>&1; tail -2 gpurun_out/oct4/ncu_crit.log
ncu -i /tmp/leaf_prof.ncu-rep --page raw --csv > gpurun_out/oct4/leaf_field_raw.csv 2>/dev/null
ncu -i /tmp/crit_prof.ncu-rep --page raw --csv > gpurun_out/oct4/criterion_raw.csv 2>/dev/null
ncu -i /tmp/leaf_prof.ncu-rep --page details > gpurun_out/oct4/leaf_field_details.txt 2>/dev/null
ncu -i /tmp/crit_prof.ncu-rep --page details > gpurun_out/oct4/criterion_details.txt 2>/dev/null
ls -la gpurun_out/oct4
