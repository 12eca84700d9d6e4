mkdir -p gpurun_out/c8
timeout 600 python -m pytest tests/test_gpu_thread.py tests/test_gpu_host_pipeline.py -x -q > gpurun_out/c8/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c8/pytest.log
tail -4 gpurun_out/c8/pytest.log
timeout 300 python bench.py --steps 20 --no-e2e 2>&1 | tail -1 | cut -c1-300
