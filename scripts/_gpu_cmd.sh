python -c "import __graft_entry__ as g; g.smoke()" > /tmp/smoke.log 2>&1; echo "smoke exit $?"; tail -2 /tmp/smoke.log
NCU=0 bash scripts/gpu_round.sh v18
cp /tmp/smoke.log gpurun_out/v18/smoke.log
grep -E "passed|failed" gpurun_out/v18/pytest_gpu.log | tail -2
