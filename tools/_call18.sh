for lag in 1 2; do echo "lag=$lag"; MP4J_SEG_LAG=$lag timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2975$lag bench/thread_coll_bench.py 2>&1 | grep collective | cut -c29-110; done
timeout 300 python -m pytest tests/test_gpu_thread_seg.py -x -q 2>&1 | tail -2
