mkdir -p gpurun_out/xc3
export MRISIM_XCHECK_OUT=$PWD/gpurun_out/xc3
(time timeout 1500 python -m pytest tests/test_ref_gpu_crosscheck.py -m gpu -x -q -rP -k cfg3 2>&1 | grep -v "^$" | tail -30) > gpurun_out/xc3/pytest.log 2>&1; grep -E "reference GPU code|passed|failed|Error|assert" gpurun_out/xc3/pytest.log | cut -c1-400; tail -4 gpurun_out/xc3/pytest.log
python -m pytest tests/test_scene_parity.py -x -q 2>&1 | tail -2
