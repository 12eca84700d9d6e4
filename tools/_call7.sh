mkdir -p gpurun_out/c7
nvidia-smi topo -m > gpurun_out/c7/topo.txt 2>&1; (lscpu | head -25; nproc; free -g; cat /proc/self/status | grep -i allowed_list) >> gpurun_out/c7/topo.txt 2>&1
for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor)" = "0x10de" ]; then echo "$d $(cat $d/class) numa=$(cat $d/numa_node) cpus=$(cat $d/local_cpulist)"; fi; done >> gpurun_out/c7/topo.txt 2>&1
CFG="1:0:64:4:-1 3:0:64:4:1:4 3:0:64:4:0:4 3:0:48:4:1:4 3:0:32:4:1:2 3:0:64:3:1:4 3:0:64:6:1:4 3:0:128:3:1:4 3:128:64:4:1:4 1:128:64:4:1 3:4:64:4:1:4 1:4:64:4:1"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 tools/ws_sweep.py --configs "$CFG" --out gpurun_out/c7/sweep8.jsonl > gpurun_out/c7/sweep8.log 2>&1
echo "rc=$?" >> gpurun_out/c7/sweep8.log
grep -v "^W\|Warning\|warn\|^\*\|OMP" gpurun_out/c7/sweep8.log | cut -c1-200 | grep -v "^--- stage trace of rank [2-7]" | head -80
CFG4="1:0:64:4:-1 3:0:64:4:1:4 3:0:64:4:0:4 3:0:48:4:1:4 3:0:32:4:1:2"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29821 tools/ws_sweep.py --configs "$CFG4" --out gpurun_out/c7/sweep4.jsonl > gpurun_out/c7/sweep4.log 2>&1
grep '"cfg"' gpurun_out/c7/sweep4.log | cut -c1-200
MP4J_WS_VER=3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29841 bench.py --gpus 8 --steps 20 > gpurun_out/c7/bench8.log 2>&1; tail -1 gpurun_out/c7/bench8.log | cut -c1-1800
MP4J_WS_VER=3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29851 bench.py --gpus 8 --steps 5 --e2e-fanout > gpurun_out/c7/bench8f.log 2>&1; tail -1 gpurun_out/c7/bench8f.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("fanout e2e", d["e2e"])'
