timeout 300 python bench/thread_coll_bench.py 2>&1 | grep collective | cut -c1-200
timeout 600 python -m pytest tests/test_gpu_thread_seg.py -x -q 2>&1 | tail -2
