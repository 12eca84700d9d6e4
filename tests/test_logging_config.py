"""config/logging.properties is a real component (role of the reference's log4j_master/log4j_slave property
files): thresholds, a separate errors-only file, per-slave local files, read by the native core."""
import os
import subprocess
import sys


def test_levels_files_and_slave_logs(tmp_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    (tmp_path / "config").mkdir()
    (tmp_path / "config" / "logging.properties").write_text(
        "master.log_dir=mlog\nmaster.file_level=INFO\nmaster.error_file=true\nmaster.echo_levels=ERROR\n"
        "slave.log_dir=slog\nslave.file_level=DEBUG\nslave.error_file=true\nslave.echo_levels=\n")
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    env.pop("MP4J_LOG_CONFIG", None)
    cmd = [sys.executable, "-m", "ytk_mp4j_b200.parallel.launch", "--nproc", "2", "--log-dir", "", "--timeout", "120",
           os.path.join(root, "tests", "log_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=200, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    mlog = (tmp_path / "mlog" / "master.log").read_text()
    merr = (tmp_path / "mlog" / "master_error.log").read_text()
    # master file threshold INFO: info + error lines of both slaves, no debug line; timestamps and levels
    assert mlog.count("an-info-line") == 2 and mlog.count("an-error-line") == 2 and "a-debug-line" not in mlog
    assert " INFO - " in mlog and "ERROR - " in mlog and "[rank=1]" in mlog
    assert merr.count("an-error-line") == 2 and "an-info-line" not in merr
    # echo_levels=ERROR: the master echoes errors only
    assert "an-error-line" in r.stderr and "an-info-line" not in r.stderr
    # every slave keeps its own files; slave threshold DEBUG
    for rank in (0, 1):
        s = (tmp_path / "slog" / ("slave_%d.log" % rank)).read_text()
        assert "a-debug-line" in s and "an-info-line" in s and "an-error-line" in s
        e = (tmp_path / "slog" / ("slave_%d_error.log" % rank)).read_text()
        assert "an-error-line" in e and "an-info-line" not in e
