"""CommCheckTool: the slave-side self-test driver (parity: J/check/CommCheckTool.java:40-105).

  python -m ytk_mp4j_b200.check.comm_check_tool <login> <masterHost> <masterPort> <arrSize> <objSize>
         <runTime> <threadNum> <mode: process|thread> <compress: true|false> <testRpc: true|false>
         [device: cpu|cuda]

Same ten positional arguments as the reference (defaults of bin/comm_cluster_error_check.sh:
arr_size=10000000 obj_size=1000 run_time=3 mode=process compress=false test_rpc=false), plus an
optional device. On any mismatch the failing slave logs it to the master and closes with code 1,
which ends the whole job (SRV:252-260); otherwise it closes with 0.
"""
import os
import sys
import threading
import traceback

from .. import ProcessCommSlave, ThreadCommSlave
from . import checks


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 10:
        print(__doc__, file=sys.stderr)
        return 64
    login, host, port = argv[0], argv[1], int(argv[2])
    arr_size, obj_size, run_time, thread_num = int(argv[3]), int(argv[4]), int(argv[5]), int(argv[6])
    mode, compress, test_rpc = argv[7], argv[8].lower() == "true", argv[9].lower() == "true"
    device = argv[10] if len(argv) > 10 else "cpu"
    if mode == "process":
        comm = ProcessCommSlave(login, host, port)
        try:
            if device == "cuda":
                comm.init_gpu()
            checks.run_all(comm, None, arr_size, obj_size, run_time, compress, test_rpc, device)
            comm.info("all checks passed")
            comm.close(0)
            return 0
        except checks.CheckFailed:
            comm.close(1)
            return 1
        except Exception as e:  # PCS:360 semantics
            comm.exception(e, sleep_s=0.5)
            return 1
    comm = ThreadCommSlave(login, thread_num, host, port)
    if device == "cuda":
        comm.init_gpu()
    failed = []

    def run(t):
        try:
            comm.setThreadId(t)
            checks.run_all(comm, t, arr_size, obj_size, run_time, compress, test_rpc, device)
        except BaseException as e:
            failed.append((e, traceback.format_exc()))
    ths = [threading.Thread(target=run, args=(t,)) for t in range(thread_num)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    if failed:
        try:
            comm.error(failed[0][1])
        finally:
            comm.close(1)
        return 1
    comm.info("all checks passed")
    comm.close(0)
    return 0


if __name__ == "__main__":
    sys.exit(main())
