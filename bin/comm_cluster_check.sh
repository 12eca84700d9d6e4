#!/bin/sh
# Cluster self-check (role of the reference's bin/comm_cluster_error_check.sh): start the master
# here, one CommCheckTool per GPU on this node. Edit the variables, then run.
slave_num=${SLAVE_NUM:-8}
thread_num=${THREAD_NUM:-1}
arr_size=${ARR_SIZE:-10000000}
obj_size=${OBJ_SIZE:-1000}
run_time=${RUN_TIME:-3}
mode=${MODE:-process}          # process | thread
compress=${COMPRESS:-false}
test_rpc=${TEST_RPC:-false}
device=${DEVICE:-cuda}         # cpu | cuda
HERE=$(cd "$(dirname "$0")/.." && pwd)
PYTHONPATH="$HERE:$PYTHONPATH" exec python -m ytk_mp4j_b200.parallel.launch --nproc "$slave_num" --threads "$thread_num" \
  --check arr_size=$arr_size obj_size=$obj_size run_time=$run_time mode=$mode compress=$compress \
  test_rpc=$test_rpc device=$device threads=$thread_num
