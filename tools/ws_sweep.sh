#!/bin/bash
# Sweeps the warp split / slot size / ring depth of the fused thread kernel through bench.py.
# usage: tools/ws_sweep.sh N "split:slotbytes:slots ..."
N=$1; shift
port=29700
for cfg in $1; do
  IFS=: read sp sb ns <<< "$cfg"
  port=$((port + 1))
  echo -n "N=$N split=$sp slot=$sb slots=$ns: "
  MP4J_WS_SPLIT=$sp MP4J_THREAD_SLOT_BYTES=$sb MP4J_THREAD_SLOTS=$ns timeout 120 python -m torch.distributed.run --nnodes=1 \
    --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2>/dev/null \
    | grep '"metric"' | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["config"]["kernel"])'
done
