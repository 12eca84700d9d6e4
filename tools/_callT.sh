mkdir -p gpurun_out/ct
timeout 900 python -m pytest tests/test_gpu_thread_seg.py -x -q > gpurun_out/ct/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ct/pytest.log
tail -8 gpurun_out/ct/pytest.log | cut -c1-300
