from .model import load_parameters  # noqa: F401
from .model_cache import ModelCache  # noqa: F401
