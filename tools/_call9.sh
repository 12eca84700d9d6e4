mkdir -p gpurun_out/c9
CFG="1:0:64:4:-1 4:0:64:4:1:4 4:0:64:4:1:1 4:0:48:4:1:4 4:0:32:4:1:2 4:0:32:6:1:2 4:0:64:3:1:4 4:0:96:3:1:4 4:0:64:4:0:4 4:128:64:4:1:4 4:4:64:4:1:4 4:1:64:4:1:4"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 tools/ws_sweep.py --configs "$CFG" --out gpurun_out/c9/sweep8.jsonl > gpurun_out/c9/sweep8.log 2>&1
echo "rc=$?" >> gpurun_out/c9/sweep8.log
grep '"cfg"\|rc=' gpurun_out/c9/sweep8.log | cut -c1-160
awk '/stage trace of rank 0/{f=1} /stage trace of rank 1/{f=0} f' gpurun_out/c9/sweep8.log | head -14
CFG4="1:0:64:4:-1 4:0:64:4:1:4 4:0:48:4:1:4 4:0:32:4:1:2 4:0:64:4:0:4 3:0:64:4:0:4"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29821 tools/ws_sweep.py --configs "$CFG4" --out gpurun_out/c9/sweep4.jsonl > gpurun_out/c9/sweep4.log 2>&1
grep '"cfg"' gpurun_out/c9/sweep4.log | cut -c1-160
for v in 1 4; do MP4J_WS_VER=$v timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2984$v bench.py --gpus 8 --steps 20 --no-e2e 2>/dev/null | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("bench N=8", d["ms_per_step"], d["config"]["kernel"], d["clocks"])'; done
