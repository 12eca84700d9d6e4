mkdir -p gpurun_out/c12
timeout 900 python -m pytest tests/test_gpu_multiproc.py -x -q > gpurun_out/c12/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c12/pytest.log
tail -12 gpurun_out/c12/pytest.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 20 > gpurun_out/c12/bench2.log 2>gpurun_out/c12/bench2.err; tail -1 gpurun_out/c12/bench2.log | cut -c1-6000; tail -5 gpurun_out/c12/bench2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 bench/thread_coll_bench.py --out gpurun_out/c12/tcb2.jsonl 2>&1 | grep collective | cut -c1-230
