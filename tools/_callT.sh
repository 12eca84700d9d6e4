timeout 900 python -m pytest tests/test_gpu_thread_seg.py -x -q 2>&1 | tail -3
timeout 300 python bench/thread_coll_bench.py 2>&1 | grep collective | cut -c1-140
