mkdir -p gpurun_out/c6
CFG="1:0:64:4:0 3:0:64:4:0:1 3:0:64:4:0:4 3:0:64:4:0:8 3:0:48:4:0:4 3:0:64:3:0:4 3:0:32:4:0:2 3:0:64:4:1:4 3:0:96:3:0:4 3:128:64:4:0:4 1:0:64:4:0 3:0:64:4:0:4 2:0:64:4:0"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 tools/ws_sweep.py --configs "$CFG" --out gpurun_out/c6/sweep2.jsonl > gpurun_out/c6/sweep2.log 2>&1
echo "rc=$?" >> gpurun_out/c6/sweep2.log
grep -v "^W\|Warning\|warn\|^\*\|OMP" gpurun_out/c6/sweep2.log | cut -c1-200
for i in 1 2 3; do timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2973$i bench.py --gpus 2 --steps 20 --no-e2e 2>/dev/null | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("bench N=2 run", d["ms_per_step"], d["config"]["kernel"])'; done
MP4J_WS_VER=3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 20 2>/dev/null | tail -1 | cut -c1-1800
