"""Slave script of tests/test_logging_config.py: logs one line per level and closes."""
import os
import sys

from ytk_mp4j_b200 import ProcessCommSlave

comm = ProcessCommSlave("logtest", os.environ["MP4J_MASTER_HOST"], int(os.environ["MP4J_MASTER_PORT"]),
                        rank_hint=int(os.environ["MP4J_RANK"]), host="127.0.0.1")
comm.debug("a-debug-line", False)
comm.info("an-info-line", False)
comm.error("an-error-line")
comm.barrier()
comm.close(0)
sys.exit(0)
