from . import conv, dense, inits  # noqa: F401
