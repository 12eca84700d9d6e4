"""CommMaster: coordination + log aggregation service (parity: J/comm/CommMaster.java:53-132,
J/rpc/Server.java). The server itself is C++ (csrc/control/master.cc); this class embeds it in
the current process. `python -m ytk_mp4j_b200.master <slave_num> <port>` and the `mp4j-master`
binary give the reference's stand-alone CLI."""
import ctypes as C
import sys

from . import _native as N
from .exception import Mp4jException, check


class CommMaster:
    def __init__(self, slave_num, port=0, connect_timeout_s=0, heartbeat_timeout_s=0,
                 exchange_timeout_s=0, watchdog_period_ms=0, log_dir=None, work_dir=None, quiet=False):
        self._h = C.c_void_p()
        check(N.lib.mp4j_master_create(int(slave_num), int(port), int(connect_timeout_s),
                                       int(heartbeat_timeout_s), int(exchange_timeout_s),
                                       int(watchdog_period_ms),
                                       log_dir.encode() if log_dir else None,
                                       work_dir.encode() if work_dir else None,
                                       1 if quiet else 0, C.byref(self._h)))
        self.slave_num = slave_num

    def start(self):
        check(N.lib.mp4j_master_start(self._h))
        return self

    @property
    def port(self):
        return N.lib.mp4j_master_port(self._h)

    def getHostPort(self):
        """CommMaster.getHostPort (J/comm/CommMaster.java): the port slaves connect to."""
        return self.port

    def getHostName(self):
        """CommMaster.getHostName: the name slaves should use to reach this master."""
        import socket
        return socket.gethostname()

    main = staticmethod(lambda argv=None: main(argv))  # CommMaster.main <slave_num> <port> (CommMaster.java:106-132)

    def stop(self, timeout_s=None):
        """Blocks until all slaves closed (CommMaster.stop, CommMaster.java:78-86); returns the
        job exit code (0 iff every slave closed with 0). With a timeout returns None if the job is
        still running."""
        code = C.c_int(0)
        rc = N.lib.mp4j_master_wait(self._h, -1 if timeout_s is None else int(timeout_s * 1000), C.byref(code))
        if rc < 0:
            raise Mp4jException(N.last_error())
        if rc == 1:
            return None
        return code.value

    def shutdown(self):
        if self._h:
            check(N.lib.mp4j_master_shutdown(self._h))

    def logs(self):
        ptr, n = C.c_void_p(), C.c_size_t()
        check(N.lib.mp4j_master_logs(self._h, C.byref(ptr), C.byref(n)))
        return N.take_bytes(ptr, n).decode("utf-8", "replace").splitlines()

    def close(self):
        if self._h:
            N.lib.mp4j_master_shutdown(self._h)
            N.lib.mp4j_master_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.close()


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 2:
        print("usage: python -m ytk_mp4j_b200.master <slave_num> <port>", file=sys.stderr)
        return 64
    m = CommMaster(int(argv[0]), int(argv[1]), log_dir="log").start()
    code = m.stop()
    m.close()
    return code


if __name__ == "__main__":
    sys.exit(main())
