from .aggregation_algorithm import AggregationAlgorithm  # noqa: F401
from .fed_avg_algorithm import FedAVGAlgorithm  # noqa: F401
from .graph_algorithm import GraphNodeEmbeddingPassingAlgorithm  # noqa: F401
from .random_dropout_algorithm import RandomDropoutAlgorithm  # noqa: F401
