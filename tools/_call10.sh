mkdir -p gpurun_out/c10
port=29840
for mode in nosampler sampler nosampler sampler; do
  port=$((port+1))
  if [ $mode = nosampler ]; then export MP4J_BENCH_NOSAMPLER=1; else unset MP4J_BENCH_NOSAMPLER; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 4 --steps 20 --no-e2e 2>gpurun_out/c10/err_$mode.log | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("bench N=4 '$mode'", d["ms_per_step"], d["config"]["kernel"], d["clocks"])'
done
unset MP4J_BENCH_NOSAMPLER
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29821 tools/ws_sweep.py --steps 20 --configs "1:0:64:4:-1 1:0:64:4:-1 3:0:64:4:0:4" 2>/dev/null | grep '"cfg"' | cut -c1-150
tail -3 gpurun_out/c10/err_sampler.log
