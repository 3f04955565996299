from .misc import accuracy  # noqa: F401
