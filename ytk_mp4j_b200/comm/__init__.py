from .process import ProcessCommSlave  # noqa: F401
from .thread import ThreadCommSlave  # noqa: F401
