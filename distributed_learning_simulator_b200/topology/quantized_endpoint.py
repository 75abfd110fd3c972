"""Endpoints that compress parameter payloads on the wire.

Same classes and switches as /root/reference/simulation_lib/topology/quantized_endpoint.py:14-116
(``Quant{Client,Server}Endpoint``, ``StochasticQuant*Endpoint``, ``NNADQ*Endpoint``, ``dequant_server_data``,
``quant_broadcast``, ``set_quant_callback``, the compression-ratio log lines).  The implementation is organised
around one idea: a message has at most ONE tensor payload field (``parameter`` or ``delta_parameter``) and a
``PayloadCodec`` packs / unpacks exactly that field, so every kind of parameter message is compressed -- the
reference only looks at ``ParameterMessage.parameter`` when sending (its default delta uploads travel
uncompressed: SURVEY.md defect D2) and asserts on anything else when receiving (D3).

The packed payload is one ``QuantizedFlat`` over the flat parameter vector; a server endpoint may leave uploads
packed (``lazy_dequant``) so that the fused aggregation kernel decodes them in registers at load time.
"""
import copy
from typing import Any, Callable

from ..algorithm.quantization import (NNADQ, NeuralNetworkAdaptiveDeterministicDequant,
                                      NeuralNetworkAdaptiveDeterministicQuant, stochastic_quantization)
from ..engine.log import get_logger
from ..message import DeltaParameterMessage, ParameterMessage
from .endpoint import ClientEndpoint, ServerEndpoint


def is_quantized(payload: Any) -> bool:
    return isinstance(payload, dict) and "__quantized__" in payload


def payload_field(message: Any) -> str | None:
    """Name of the tensor-payload attribute of ``message`` (None: nothing to compress)."""
    if isinstance(message, DeltaParameterMessage):
        return "delta_parameter"
    if isinstance(message, ParameterMessage):
        return "parameter"
    return None


class PayloadCodec:
    """A (pack, unpack) pair applied to the payload field of parameter messages; other objects pass through."""

    def __init__(self, pack: Callable | None, unpack: Callable) -> None:
        self.pack, self.unpack = pack, unpack

    def _apply(self, message: Any, fn: Callable | None) -> bool:
        field = payload_field(message)
        if field is None or fn is None:
            return False
        setattr(message, field, fn(getattr(message, field)))
        return True

    def encode(self, message: Any) -> bool:
        return self._apply(message, self.pack)

    def decode(self, message: Any) -> bool:
        return self._apply(message, self.unpack)


class QuantClientEndpoint(ClientEndpoint):
    def __init__(self, quant: Callable, dequant: Callable, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._codec = PayloadCodec(quant, dequant)
        self.dequant_server_data: bool = False

    def get(self) -> Any:
        message = super().get()
        if self.dequant_server_data:
            self._codec.decode(message)
        return message

    def send(self, data: Any) -> None:
        if self._codec.encode(data):
            self._after_quant(data=data)
        super().send(data=data)

    def _after_quant(self, data: Any) -> None:
        """Hook: ``data`` now carries the packed payload."""


class QuantServerEndpoint(ServerEndpoint):
    def __init__(self, quant: Callable | None, dequant: Callable, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._codec = PayloadCodec(quant, dequant)
        self.quant_broadcast: bool = False
        self.lazy_dequant: bool = False  # leave uploads packed for the fused reduce kernel

    def get(self, worker_id: int) -> Any:
        message = super().get(worker_id=worker_id)
        if not self.lazy_dequant:
            self._codec.decode(message)
        return message

    def dequant(self, payload: Any) -> Any:
        return self._codec.unpack(payload)

    def broadcast(self, data: Any, **kwargs: Any) -> None:
        if self.quant_broadcast and payload_field(data) is not None:
            if self._codec.pack is None:
                raise RuntimeError("quant_broadcast needs a quantizer on the server endpoint")
            # the server keeps its fp32 model: pack a shallow copy for the wire
            wire = copy.copy(data)
            wire.other_data = dict(data.other_data)
            self._codec.encode(wire)
            self._after_quant(data=wire)
            data = wire
        return super().broadcast(data=data, **kwargs)

    def _after_quant(self, data: Any) -> None:
        """Hook: ``data`` is the packed copy about to be broadcast."""


class StochasticQuantClientEndpoint(QuantClientEndpoint):
    def __init__(self, **kwargs: Any) -> None:
        quant, dequant = stochastic_quantization(quantization_level=255)
        super().__init__(quant=quant, dequant=dequant, **kwargs)


class StochasticQuantServerEndpoint(QuantServerEndpoint):
    def __init__(self, **kwargs: Any) -> None:
        quant, dequant = stochastic_quantization(quantization_level=255)
        super().__init__(quant=quant, dequant=dequant, **kwargs)


def _log_ratio(data: Any, prefix: str) -> None:
    # analysis/analyze_log.py scrapes "<prefix> NNABQ compression ratio is <float>"
    NeuralNetworkAdaptiveDeterministicQuant.check_compression_ratio(quantized_data=data, prefix=prefix)


class NNADQClientEndpoint(QuantClientEndpoint):
    def __init__(self, weight: float, **kwargs: Any) -> None:
        get_logger().info("use weight %s", weight)
        quant, dequant = NNADQ(weight=weight)
        super().__init__(quant=quant, dequant=dequant, **kwargs)

    def _after_quant(self, data: Any) -> None:
        _log_ratio(data, "worker")


class NNADQServerEndpoint(QuantServerEndpoint):
    """``weight=None``: a dequantize-only server (it never quantizes its broadcasts)."""

    def __init__(self, weight: float | None = None, **kwargs: Any) -> None:
        quant, dequant = (None, NeuralNetworkAdaptiveDeterministicDequant()) if weight is None else NNADQ(weight=weight)
        super().__init__(quant=quant, dequant=dequant, **kwargs)
        self.__quant_callback: Callable | None = None

    def set_quant_callback(self, quant_callback: Callable) -> None:
        self.__quant_callback = quant_callback

    def _after_quant(self, data: Any) -> None:
        _log_ratio(data, "broadcast")
        if self.__quant_callback is not None:
            self.__quant_callback(data=data)
