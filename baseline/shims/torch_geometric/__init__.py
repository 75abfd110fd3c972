from . import nn, utils  # noqa: F401
