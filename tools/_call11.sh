mkdir -p gpurun_out/c11
CFG="1:0:64:4:-1:4:0 1:0:64:4:-1:4:1 1:0:64:3:-1:4:1 1:0:48:4:-1:4:1 1:0:96:3:-1:4:1 1:0:32:6:-1:4:1 1:128:64:4:-1:4:1 1:4:64:4:-1:4:1 1:0:64:4:-1:4:0 1:0:64:4:-1:4:1"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 tools/ws_sweep.py --configs "$CFG" --out gpurun_out/c11/sweep8.jsonl > gpurun_out/c11/sweep8.log 2>&1
echo "rc=$?" >> gpurun_out/c11/sweep8.log
grep '"cfg"\|rc=' gpurun_out/c11/sweep8.log | cut -c1-150
awk '/stage trace of rank 0/{f=1} /stage trace of rank 1/{f=0} f' gpurun_out/c11/sweep8.log | head -14
CFG4="1:0:64:4:-1:4:0 1:0:64:4:-1:4:1 1:0:48:4:-1:4:1 1:0:32:4:-1:4:1 3:0:32:4:0:2"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29821 tools/ws_sweep.py --configs "$CFG4" --out gpurun_out/c11/sweep4.jsonl 2>/dev/null | grep '"cfg"' | cut -c1-150
CFG2="1:0:64:4:0:4:0 1:0:64:4:0:4:1 1:0:32:4:0:4:1 1:0:48:4:0:4:1 3:0:32:4:0:2 3:0:32:4:0:4 3:0:32:5:0:2 3:0:24:4:0:2"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29831 tools/ws_sweep.py --configs "$CFG2" --out gpurun_out/c11/sweep2.jsonl 2>/dev/null | grep '"cfg"' | cut -c1-150
