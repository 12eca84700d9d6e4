mkdir -p gpurun_out/c16
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench/codec_bench.py 2>&1 | grep allreduce_wire | tee gpurun_out/c16/codec2.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 bench/thread_coll_bench.py --out gpurun_out/c16/tcb2.jsonl 2>&1 | grep collective | cut -c1-170
timeout 600 python -m pytest tests/test_gpu_multiproc.py -x -q 2>&1 | tail -3
