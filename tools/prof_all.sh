#!/bin/bash
# One `ncu --set full` capture per kernel family on ONE GPU (emulated ranks), after a plain run of the same
# command exited 0. Reports go to gpurun_out/ncu/<name>.ncu-rep; read them with tools/ncu_summary.py.
# Usage: tools/prof_all.sh [name ...]   (no names = every family)
set -u
mkdir -p gpurun_out/ncu
only=" $* "
run() {  # name, --which, kernel regex, launches to skip
  local name=$1 which=$2 regex=$3 skip=$4
  if [ "$only" != "  " ] && [[ "$only" != *" $name "* ]]; then return; fi
  python tools/prof_all.py --which $which --iters 3 > gpurun_out/ncu/$name.plain.log 2>&1 || { echo "$name: plain run failed"; tail -3 gpurun_out/ncu/$name.plain.log; return; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 \
      -o gpurun_out/ncu/$name -f python tools/prof_all.py --which $which --iters 3 > gpurun_out/ncu/$name.ncu.log 2>&1
  echo "$name: rc=$? $(ls -la gpurun_out/ncu/$name.ncu-rep 2>/dev/null | awk '{print $5}') bytes"
  # the reports (with imported source) are tens of MB each and gpurun copies back 64 MiB at most: condense them
  # here into the raw metric page and the per-line source page, keep the report itself only if it is small
  ncu -i gpurun_out/ncu/$name.ncu-rep --page raw --csv > gpurun_out/ncu/$name.raw.csv 2>/dev/null
  ncu -i gpurun_out/ncu/$name.ncu-rep --page source --csv 2>/dev/null | head -c 3000000 > gpurun_out/ncu/$name.source.csv
  if [ "$(stat -c %s gpurun_out/ncu/$name.ncu-rep 2>/dev/null || echo 0)" -gt 6000000 ]; then rm -f gpurun_out/ncu/$name.ncu-rep; fi
}
run grs_twoshot grs_twoshot grs_kernel 1
run grs_oneshot grs_oneshot grs_kernel 1
run thread_ws ws thread_fused_ws_kernel 1
run seg_rs seg_rs thread_seg_kernel 1
run seg_ag seg_ag thread_seg_kernel 1
run map_partition map map_partition_push_kernel 1
run map_table map map_table_kernel 1
run map_gather map map_allgather_push_kernel 1
run gbdt gbdt gbdt_hist 1
run codec codec codec_allreduce_kernel 1
run lr_xtr lr lr_xtr 1
run lr_residual lr_residual lr_residual_kernel 1
grep -h "iter 2" gpurun_out/ncu/*.plain.log
