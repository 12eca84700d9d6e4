"""Native (C++) multi-rank stress test of the host plane, plain and under ThreadSanitizer /
AddressSanitizer (the reference has no race detection at all). One process, p ranks as
threads, every dense collective x 6 dtypes + operators + mailbox + byte transport + exchange."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import build as mp4j_build  # noqa: E402


def _run(exe, *args, shm="1"):
    # the TCP ring sends big steps on a helper thread: a 16 KiB threshold puts that path under the sanitizers too
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66", ASAN_OPTIONS="detect_leaks=1", MP4J_SHM=shm,
               MP4J_SHM_BYTES=str(1 << 20), MP4J_TCP_DUPLEX_MIN_BYTES=str(16 << 10))
    import tempfile
    with tempfile.TemporaryDirectory(prefix="mp4j-") as cwd:  # the master drops kill_<port>.sh into its cwd
        return subprocess.run([exe] + [str(a) for a in args], capture_output=True, text=True, timeout=300, env=env, cwd=cwd)


@pytest.mark.parametrize("p,rounds,n", [(1, 1, 1000), (2, 2, 100003), (3, 2, 1000), (4, 2, 20011), (8, 2, 5000)])
@pytest.mark.parametrize("shm", ["1", "0"])
def test_host_stress_plain(p, rounds, n, shm):
    exe = mp4j_build.build_cpp_tests(None)
    r = _run(exe, p, rounds, n, shm=shm)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ok (0 failures)" in r.stdout


@pytest.mark.parametrize("sanitizer", ["thread", "address"])
def test_host_stress_sanitized(sanitizer):
    try:
        exe = mp4j_build.build_cpp_tests(sanitizer)
    except RuntimeError as e:  # no sanitizer runtime on this machine
        pytest.skip("cannot build with -fsanitize=%s: %s" % (sanitizer, str(e)[-200:]))
    for p, rounds, n, shm in ((4, 2, 20011, "1"), (8, 2, 3000, "1"), (4, 2, 300011, "1"), (4, 2, 20011, "0")):
        r = _run(exe, p, rounds, n, shm=shm)
        report = r.stdout + r.stderr
        assert "WARNING: ThreadSanitizer" not in report and "ERROR: AddressSanitizer" not in report, report[-4000:]
        assert "ERROR: LeakSanitizer" not in report, report[-4000:]
        assert r.returncode == 0, report[-2000:]
