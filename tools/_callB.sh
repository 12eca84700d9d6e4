mkdir -p gpurun_out/cb
timeout 300 python bench/thread_coll_bench.py --out gpurun_out/cb/tcb1.jsonl 2>&1 | grep collective | cut -c1-260
