"""Local cluster harness: master + N slave processes on this host.

The reference has no in-process fake transport or loopback harness; its only
multi-node story is ssh + scp (bin/comm_cluster_error_check.sh). This module is what the tests,
the check tool and `mp4j-launch` use to run N ranks (optionally x T threads) locally.
"""
import multiprocessing as mp
import os
import tempfile
import threading
import traceback


def _worker(fn, rank_hint, n, port, threads, args, q, env):
    os.environ.update(env or {})
    # tests of the multi-machine paths on one box: MP4J_NODE_IDS="0,0,1,1" gives rank i the machine id of entry i
    if env and env.get("MP4J_NODE_IDS") and rank_hint is not None:
        os.environ["MP4J_NODE_ID"] = env["MP4J_NODE_IDS"].split(",")[rank_hint]
    try:
        from ytk_mp4j_b200 import ProcessCommSlave, ThreadCommSlave
        if not threads or threads < 1:
            comm = ProcessCommSlave("test", "127.0.0.1", port, rank_hint=rank_hint, host="127.0.0.1")
            try:
                out = fn(comm, *args)
                comm.close(0)
                q.put((rank_hint, comm.getRank(), "ok", out))
            except BaseException:
                tb = traceback.format_exc()
                try:
                    comm.error(tb)
                    comm.close(1)
                except Exception:
                    pass
                q.put((rank_hint, -1, "err", tb))
        else:
            comm = ThreadCommSlave("test", threads, "127.0.0.1", port, rank_hint=rank_hint, host="127.0.0.1")
            outs, errs = [None] * threads, []

            def run(t):
                try:
                    comm.setThreadId(t)
                    outs[t] = fn(comm, t, *args)
                except BaseException:
                    errs.append(traceback.format_exc())
            ths = [threading.Thread(target=run, args=(t,)) for t in range(threads)]
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            if errs:
                try:
                    comm.error(errs[0])
                    comm.close(1)
                except Exception:
                    pass
                q.put((rank_hint, -1, "err", errs[0]))
            else:
                comm.close(0)
                q.put((rank_hint, comm.getRank(), "ok", outs))
    except BaseException:
        q.put((rank_hint, -1, "err", traceback.format_exc()))


def run_cluster(n, fn, args=(), threads=0, timeout=120, use_hints=True, env=None, master_kwargs=None):
    """Runs fn(comm, *args) (or fn(comm, tid, *args) on a ThreadCommSlave when threads >= 1) on n local slave
    processes against an embedded master. Returns (results_by_rank, master_exit_code, master_logs).
    Raises RuntimeError with the first traceback if any slave failed."""
    from ytk_mp4j_b200 import CommMaster
    ctx = mp.get_context("spawn")
    mk = dict(master_kwargs or {})
    mk.setdefault("work_dir", tempfile.mkdtemp(prefix="mp4j-"))  # kill_<port>.sh and writeFile outputs
    master = CommMaster(n, 0, quiet=True, **mk).start()
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(fn, i if use_hints else None, n, master.port, threads, args, q, env))
             for i in range(n)]
    for p in procs:
        p.start()
    results, errors = {}, []
    try:
        for _ in range(n):
            hint, rank, status, out = q.get(timeout=timeout)
            if status == "ok":
                results[rank] = out
            else:
                errors.append(out)
    except Exception as e:  # queue.Empty
        errors.append("timeout waiting for slaves: %r" % (e,))
    for p in procs:
        p.join(timeout=10)
        if p.is_alive():
            p.kill()
    code = master.stop(timeout_s=5)
    logs = master.logs()
    master.close()
    if errors:
        raise RuntimeError("slave failure:\n" + errors[0])
    return [results[r] for r in range(n)], code, logs
