from .executor import Trainer  # noqa: F401
