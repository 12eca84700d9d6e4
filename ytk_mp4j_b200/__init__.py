"""mp4j-b200: a Blackwell-native collective message-passing framework with the capabilities and
public API of ytk-mp4j (CommMaster, ProcessCommSlave, ThreadCommSlave, Operands, Operators,
array + map containers), built on one C++/CUDA core (libmp4j.so, include/mp4j.h)."""
from .exception import Mp4jException  # noqa: F401
from .operators import Operators, Operator  # noqa: F401
from .operands import Operands, Operand, PickleSerializer  # noqa: F401
from .master import CommMaster  # noqa: F401
from .comm.process import ProcessCommSlave  # noqa: F401
from .comm.thread import ThreadCommSlave  # noqa: F401
from .utils.comm_utils import CommUtils  # noqa: F401
from .operator_enums import Collective, Container  # noqa: F401

__version__ = "0.1.0"
