mkdir -p gpurun_out/c13
timeout 900 python -m pytest tests/test_gpu_thread_seg.py tests/test_gpu_multiproc.py -x -q > gpurun_out/c13/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c13/pytest.log
tail -6 gpurun_out/c13/pytest.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 bench/thread_coll_bench.py --out gpurun_out/c13/tcb2.jsonl 2>&1 | grep collective | cut -c1-230
timeout 300 python bench/thread_coll_bench.py --out gpurun_out/c13/tcb1.jsonl 2>&1 | grep collective | cut -c1-230
