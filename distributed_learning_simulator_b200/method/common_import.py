"""Names every method package needs (mirrors the role of
/root/reference/simulation_lib/method/common_import.py)."""
from ..algorithm.fed_avg_algorithm import FedAVGAlgorithm  # noqa: F401
from ..message import (DeltaParameterMessage, Message, ParameterMessage, ParameterMessageBase,  # noqa: F401
                       get_message_size)
from ..server.aggregation_server import AggregationServer  # noqa: F401
from ..server.graph_server import GraphNodeServer  # noqa: F401
from ..topology.quantized_endpoint import (NNADQClientEndpoint, NNADQServerEndpoint, QuantClientEndpoint,  # noqa: F401
                                           QuantServerEndpoint, StochasticQuantClientEndpoint,
                                           StochasticQuantServerEndpoint)
from ..worker.aggregation_worker import AggregationWorker  # noqa: F401
from ..worker.error_feedback_worker import ErrorFeedbackWorker  # noqa: F401
from ..worker.gradient_worker import GradientWorker  # noqa: F401
from ..worker.graph_worker import GraphWorker  # noqa: F401
