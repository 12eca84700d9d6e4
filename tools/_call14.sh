mkdir -p gpurun_out/c14
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench/sweep.py --out gpurun_out/c14/sweep_4.jsonl > gpurun_out/c14/sweep4.log 2>&1; echo "sweep rc=$?"; tail -3 gpurun_out/c14/sweep4.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 4 --steps 20 > gpurun_out/c14/bench4.log 2>gpurun_out/c14/bench4.err; echo "bench rc=$?"
tail -1 gpurun_out/c14/bench4.log | python -c '
import sys, json
d = json.loads(sys.stdin.read())
print("ms", d["ms_per_step"], d["config"]["kernel"], "e2e", d["e2e"]["ms_per_step"], d["clocks"])
print("correctness", d["correctness_matrix"])
print(json.dumps(d["consumers"]))
sw = d["dense_sweep"]["float"]
for c in ("allreduce", "reduce_scatter", "allgather"):
    print(c, "sym", sw[c]["symmetric"][-4:], "ord", sw[c]["ordinary"][-4:], "nccl", sw[c]["nccl"][-4:], sw[c]["picked"][-4:])
'
tail -3 gpurun_out/c14/bench4.err
