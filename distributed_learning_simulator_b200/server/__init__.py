from .aggregation_server import AggregationServer  # noqa: F401
from .graph_server import GraphNodeServer  # noqa: F401
from .server import Server  # noqa: F401
