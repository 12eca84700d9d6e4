mkdir -p gpurun_out/c17
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus $N --steps 20 > gpurun_out/c17/bench$N.log 2>gpurun_out/c17/bench$N.err; echo "bench rc=$?"
tail -1 gpurun_out/c17/bench$N.log | python -c '
import sys, json
d = json.loads(sys.stdin.read())
print("ms", d["ms_per_step"], d["config"]["kernel"], "e2e", d["e2e"]["ms_per_step"], d["clocks"])
print("correctness", d["correctness_matrix"])
print(json.dumps(d["consumers"]))
for t in ("float", "double"):
    sw = d["dense_sweep"][t]
    for c in ("allreduce", "reduce_scatter", "allgather"):
        print(t, c, "sym", sw[c]["symmetric"][-4:], "ord", sw[c]["ordinary"][-4:], "nccl", sw[c]["nccl"][-4:], sw[c]["picked"][-4:])
'
tail -2 gpurun_out/c17/bench$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 bench/thread_coll_bench.py --out gpurun_out/c17/tcb$N.jsonl 2>&1 | grep collective | cut -c1-150
