mkdir -p gpurun_out/c15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 8 --steps 20 > gpurun_out/c15/bench8.log 2>gpurun_out/c15/bench8.err; echo "bench rc=$?"
tail -1 gpurun_out/c15/bench8.log | python -c '
import sys, json
d = json.loads(sys.stdin.read())
print("ms", d["ms_per_step"], d["config"]["kernel"], "e2e", d["e2e"]["ms_per_step"], d["clocks"])
print("correctness", d["correctness_matrix"])
print(json.dumps(d["consumers"]))
for t in ("float", "double"):
    sw = d["dense_sweep"][t]
    for c in ("allreduce", "reduce_scatter", "allgather"):
        print(t, c, "sym", sw[c]["symmetric"][-4:], "ord", sw[c]["ordinary"][-4:], "nccl", sw[c]["nccl"][-4:], sw[c]["picked"][-4:])
    print(t, sw["allreduce_oneshot_latency_us"])
'
tail -2 gpurun_out/c15/bench8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench/sweep.py --out gpurun_out/c15/sweep_8.jsonl > gpurun_out/c15/sweep8.log 2>&1; echo "sweep rc=$?"; tail -2 gpurun_out/c15/sweep8.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29751 bench/thread_coll_bench.py --out gpurun_out/c15/tcb8.jsonl 2>&1 | grep collective | cut -c1-200
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29771 bench/lr_bench.py 2>&1 | grep lr_xtr | cut -c1-400
