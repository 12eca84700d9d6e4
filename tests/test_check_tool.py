"""CommCheckTool equivalent: the whole check matrix through the launcher (CPU, loopback)."""
import subprocess
import sys

import pytest


@pytest.mark.parametrize("mode,threads,nproc", [("process", 1, 2), ("thread", 2, 2)])
def test_check_matrix_via_launcher(mode, threads, nproc, tmp_path):
    cmd = [sys.executable, "-m", "ytk_mp4j_b200.parallel.launch", "--nproc", str(nproc), "--threads", str(threads),
           "--log-dir", str(tmp_path / "log"), "--timeout", "240", "--check", "arr_size=2000", "obj_size=30",
           "run_time=1", "mode=%s" % mode, "threads=%d" % threads, "test_rpc=true", "compress=true"]
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    # run from the temp dir: the master drops kill_<port>.sh into its working directory, like the reference
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    log = (tmp_path / "log" / "master.log").read_text()
    assert "all checks passed" in log and "allreduce takes:" in log


def test_multi_host_launch_through_ssh_command(tmp_path):
    """--hosts: slaves started through `<ssh> <host> <command>`; here ssh is a local stand-in and both "hosts"
    are this machine, so the whole remote command line (cd, env, quoting, LOCAL_RANK per host) is exercised."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "ytk_mp4j_b200.parallel.launch", "--nproc", "4", "--hosts", "127.0.0.1,127.0.0.1",
           "--ssh", "%s %s" % (sys.executable, os.path.join(root, "tests", "fake_ssh.py")),
           "--log-dir", str(tmp_path / "log"), "--timeout", "240", "--check", "arr_size=2000", "obj_size=30",
           "run_time=1", "mode=process"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "all checks passed" in (tmp_path / "log" / "master.log").read_text()
