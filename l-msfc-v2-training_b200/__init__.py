from . import spec  # noqa: F401
