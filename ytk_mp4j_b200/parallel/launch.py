"""mp4j-launch: start a master plus N local slaves (one per GPU), optionally T threads each.

Replaces bin/comm_cluster_error_check.sh (kill the previous run, scp artefacts, start the master,
ssh-start one JVM per host; R/bin/comm_cluster_error_check.sh:1-61) : single node by default (one slave per GPU), or one slave group per host over ssh (--hosts).

  python -m ytk_mp4j_b200.parallel.launch --nproc 8 [--port 61235] [--threads 1] script.py [args...]
  python -m ytk_mp4j_b200.parallel.launch --nproc 4 --check arr_size=1000000 mode=thread threads=2 device=cuda
  python -m ytk_mp4j_b200.parallel.launch --nproc 16 --hosts node1,node2 --host node1 script.py   # 8 slaves per host

With --hosts the slaves are started through `<ssh> <host> '<command>'` (--ssh, default "ssh -o BatchMode=yes"),
in the launcher's working directory on every host (the tree must be visible there, e.g. a shared
filesystem; the reference scp's its lib/ config/ bin/ instead), --nproc / len(hosts) slaves per host with
LOCAL_RANK = index on the host. --kill-previous first runs a kill_<port>.sh left by an earlier master on
this port (the reference's script does the same with its pid file).

Each slave process gets MP4J_MASTER_HOST, MP4J_MASTER_PORT, MP4J_RANK (rank hint), MP4J_WORLD,
LOCAL_RANK and CUDA device = LOCAL_RANK. The launcher's exit code is the master's job code
(0 iff every slave closed with 0), exactly what CommMaster.main returns (CommMaster.java:127-130).
"""
import argparse
import os
import subprocess
import sys

from ..master import CommMaster


def main(argv=None):
    ap = argparse.ArgumentParser(prog="mp4j-launch")
    ap.add_argument("--nproc", type=int, required=True)
    ap.add_argument("--port", type=int, default=0)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--log-dir", default="log")
    ap.add_argument("--log-config", default=None, help="logging.properties (default: ./config/logging.properties "
                    "when present); levels, files and rolling of master and slave logs")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("--check", action="store_true", help="run the CommCheckTool matrix instead of a script")
    ap.add_argument("--print-ssh", nargs="*", metavar="HOST", help="only print per-host ssh command lines")
    ap.add_argument("--hosts", default=None, help="comma-separated hosts: start the slaves there over ssh")
    ap.add_argument("--ssh", default="ssh -o BatchMode=yes", help="remote shell command (host and command are appended)")
    ap.add_argument("--kill-previous", action="store_true", help="run kill_<port>.sh of an earlier job on this port first")
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)

    if a.kill_previous and a.port and os.path.exists("kill_%d.sh" % a.port):
        subprocess.call(["sh", "kill_%d.sh" % a.port], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    # logging configuration: read by the native core of the master (this process) and of every slave
    cfg = a.log_config or ("config/logging.properties" if os.path.exists("config/logging.properties") else None)
    if cfg and "MP4J_LOG_CONFIG" not in os.environ:
        os.environ["MP4J_LOG_CONFIG"] = os.path.abspath(cfg)
    master = CommMaster(a.nproc, a.port, log_dir=a.log_dir, quiet=False).start()
    port = master.port
    if a.check:
        kv = dict(x.split("=", 1) for x in a.rest if "=" in x)
        cmd = [sys.executable, "-m", "ytk_mp4j_b200.check.comm_check_tool", os.environ.get("USER", "mp4j"), a.host,
               str(port), kv.get("arr_size", "1000000"), kv.get("obj_size", "1000"), kv.get("run_time", "1"),
               kv.get("threads", str(a.threads)), kv.get("mode", "process"), kv.get("compress", "false"),
               kv.get("test_rpc", "false"), kv.get("device", "cpu")]
    else:
        if not a.rest:
            ap.error("give a script to run (or --check)")
        cmd = [sys.executable] + [x for x in a.rest if x != "--"]
    if a.print_ssh is not None:
        for i, h in enumerate(a.print_ssh or ["<host>"] * a.nproc):
            print("ssh %s 'MP4J_MASTER_HOST=%s MP4J_MASTER_PORT=%d MP4J_RANK=%d MP4J_WORLD=%d %s'" %
                  (h, a.host, port, i, a.nproc, " ".join(cmd)))
        master.close()
        return 0
    procs = []
    if a.hosts:
        import shlex
        hosts = [h for h in a.hosts.split(",") if h]
        if a.nproc % len(hosts) != 0:
            ap.error("--nproc must be a multiple of the number of hosts")
        per = a.nproc // len(hosts)
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        for i in range(a.nproc):
            h, local = hosts[i // per], i % per
            envs = {"MP4J_MASTER_HOST": a.host, "MP4J_MASTER_PORT": port, "MP4J_RANK": i, "MP4J_WORLD": a.nproc,
                    "LOCAL_RANK": local, "MP4J_HOST": h, "MP4J_THREADS": a.threads,
                    "PYTHONPATH": root + os.pathsep + os.environ.get("PYTHONPATH", "")}
            if os.environ.get("MP4J_LOG_CONFIG"):
                envs["MP4J_LOG_CONFIG"] = os.environ["MP4J_LOG_CONFIG"]
            remote = "cd %s && env %s %s" % (shlex.quote(os.getcwd()),
                                            " ".join("%s=%s" % (k, shlex.quote(str(v))) for k, v in envs.items()),
                                            " ".join(shlex.quote(c) for c in cmd))
            procs.append(subprocess.Popen(shlex.split(a.ssh) + [h, remote]))
    for i in range(a.nproc if not a.hosts else 0):
        env = dict(os.environ, MP4J_MASTER_HOST=a.host, MP4J_MASTER_PORT=str(port), MP4J_RANK=str(i),
                   MP4J_WORLD=str(a.nproc), LOCAL_RANK=str(i), MP4J_HOST="127.0.0.1", MP4J_THREADS=str(a.threads))
        procs.append(subprocess.Popen(cmd, env=env))
    code = master.stop(timeout_s=a.timeout)
    for p in procs:
        try:
            p.wait(timeout=30 if code is not None else 1)
        except subprocess.TimeoutExpired:
            p.kill()  # the exact children this launcher started
    master.close()
    if code is None:
        print("mp4j-launch: timeout, job killed", file=sys.stderr)
        return 124
    print("mp4j-launch: job exit code %d" % code, file=sys.stderr)
    return code


if __name__ == "__main__":
    sys.exit(main())
