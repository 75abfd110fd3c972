from .executor import Inferencer  # noqa: F401
