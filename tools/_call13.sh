mkdir -p gpurun_out/c13
timeout 900 python -m pytest tests/test_gpu_multiproc.py -x -q > gpurun_out/c13/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c13/pytest.log
tail -25 gpurun_out/c13/pytest.log | cut -c1-300
