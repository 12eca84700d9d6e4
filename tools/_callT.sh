mkdir -p gpurun_out/ct
timeout 1200 python -m pytest tests/test_gpu_multiproc.py -x -q > gpurun_out/ct/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ct/pytest.log
tail -25 gpurun_out/ct/pytest.log | cut -c1-250
