mkdir -p gpurun_out/ct
timeout 1200 python -m pytest tests/test_gpu_thread_seg.py tests/test_gpu_watchdog.py tests/test_zz_no_alloc.py -x -q > gpurun_out/ct/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ct/pytest.log
tail -25 gpurun_out/ct/pytest.log | cut -c1-220
