from .deterministic import (NNADQ, NeuralNetworkAdaptiveDeterministicDequant,  # noqa: F401
                            NeuralNetworkAdaptiveDeterministicQuant)
from .packed import QuantizedFlat  # noqa: F401
from .stochastic import stochastic_quantization  # noqa: F401
