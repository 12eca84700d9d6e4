"""Mp4jException: the one error type raised by every API (parity: J/exception/Mp4jException.java:30-44)."""


class Mp4jException(Exception):
    pass


def check(rc):
    """Raises Mp4jException with the native message if a C ABI call failed."""
    if rc != 0:
        from . import _native
        raise Mp4jException(_native.last_error())
