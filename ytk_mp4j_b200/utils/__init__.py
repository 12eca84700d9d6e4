from .comm_utils import CommUtils  # noqa: F401
from . import scatter_allocate  # noqa: F401
