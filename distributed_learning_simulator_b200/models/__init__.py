"""Model zoo: every model name that appears in the reference's ``conf/**`` tree plus
ResNet-18 (BASELINE.json config 2).  The reference resolves these names through the
external ``cyy_torch_vision/text/graph`` packages
(/root/reference/simulation_lib/method/common_import.py:1-2,13-16)."""
from .registry import create_model, get_model_names, register_model  # noqa: F401
from . import vision, text, graph  # noqa: F401  (registration side effects)
from .graph import GCNConv, MessagePassing  # noqa: F401
