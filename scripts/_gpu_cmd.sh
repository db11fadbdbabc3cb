python -c "import __graft_entry__ as g; g.smoke()" > /tmp/smoke.log 2>&1; echo "smoke exit $?"; tail -2 /tmp/smoke.log
bash scripts/gpu_round.sh v12
cp /tmp/smoke.log gpurun_out/v12/smoke.log
for wl in cfg1_default_4032_waters cfg2_8000_waters_intracellular_kick; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/v12/bench_${wl}.json 2>> gpurun_out/v12/bench.err
done
rm -f gpurun_out/v12/prof.ncu-rep.tmp; ls -la gpurun_out/v12
