from .endpoint import ClientEndpoint, ServerEndpoint  # noqa: F401
from .quantized_endpoint import (NNADQClientEndpoint, NNADQServerEndpoint, QuantClientEndpoint,  # noqa: F401
                                 QuantServerEndpoint, StochasticQuantClientEndpoint,
                                 StochasticQuantServerEndpoint)
from .topology import CentralTopology, ProcessPipeCentralTopology, ThreadCentralTopology  # noqa: F401
