from . import engine  # noqa: F401
