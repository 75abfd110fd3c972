"""Endpoints that compress parameter payloads on the wire.

Feature parity with /root/reference/simulation_lib/topology/quantized_endpoint.py:14-116
(``Quant{Client,Server}Endpoint``, ``StochasticQuant*Endpoint``, ``NNADQ*Endpoint``,
``dequant_server_data``, ``quant_broadcast``, ``set_quant_callback``, the compression
ratio log lines), with the reference's defects fixed:

* D2 -- the client also quantizes ``DeltaParameterMessage`` payloads (the default upload
  kind), so FedPAQ uploads and FedOBD stage-2 uploads really travel compressed;
* D3 -- ``get`` on the client dequantizes the payload once and passes ``None`` / plain
  ``Message`` objects through instead of asserting.

The quantized payload is one ``QuantizedFlat`` over the flat parameter vector; the server
endpoint may leave it packed (``lazy_dequant``) so the fused aggregation kernel decodes
it in registers at load time instead of materialising fp32 copies per worker.
"""
from typing import Any, Callable

from ..algorithm.quantization import (NNADQ, NeuralNetworkAdaptiveDeterministicDequant,
                                      NeuralNetworkAdaptiveDeterministicQuant, stochastic_quantization)
from ..engine.log import get_logger
from ..message import DeltaParameterMessage, ParameterMessage, ParameterMessageBase
from .endpoint import ClientEndpoint, ServerEndpoint


def is_quantized(payload: Any) -> bool:
    return isinstance(payload, dict) and "__quantized__" in payload


class QuantClientEndpoint(ClientEndpoint):
    def __init__(self, quant: Callable, dequant: Callable, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._quant: Callable = quant
        self._dequant: Callable = dequant
        self.dequant_server_data: bool = False

    def get(self) -> Any:
        data = super().get()
        if not self.dequant_server_data:
            return data
        if isinstance(data, ParameterMessage):
            data.parameter = self._dequant(data.parameter)
        elif isinstance(data, DeltaParameterMessage):
            data.delta_parameter = self._dequant(data.delta_parameter)
        return data

    def send(self, data: Any) -> None:
        if isinstance(data, ParameterMessage):
            data.parameter = self._quant(data.parameter)
            self._after_quant(data=data)
        elif isinstance(data, DeltaParameterMessage):
            data.delta_parameter = self._quant(data.delta_parameter)
            self._after_quant(data=data)
        super().send(data=data)

    def _after_quant(self, data: Any) -> None:
        pass


class QuantServerEndpoint(ServerEndpoint):
    def __init__(self, quant: Callable | None, dequant: Callable, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._quant: Callable | None = quant
        self._dequant: Callable = dequant
        self.quant_broadcast: bool = False
        self.lazy_dequant: bool = False  # leave uploads packed for the fused reduce kernel

    def get(self, worker_id: int) -> Any:
        data = super().get(worker_id=worker_id)
        if self.lazy_dequant:
            return data
        match data:
            case ParameterMessage():
                data.parameter = self._dequant(data.parameter)
            case DeltaParameterMessage():
                data.delta_parameter = self._dequant(data.delta_parameter)
        return data

    def dequant(self, payload: Any) -> Any:
        return self._dequant(payload)

    def broadcast(self, data: Any, **kwargs: Any) -> None:
        if self.quant_broadcast and isinstance(data, ParameterMessageBase):
            assert isinstance(data, ParameterMessage)
            assert self._quant is not None
            # the server keeps its own fp32 copy: quantize a shallow copy for the wire
            import copy

            wire = copy.copy(data)
            wire.other_data = dict(data.other_data)
            wire.parameter = self._quant(data.parameter)
            get_logger().debug("broadcast quantization")
            self._after_quant(data=wire)
            data = wire
        return super().broadcast(data=data, **kwargs)

    def _after_quant(self, data: Any) -> None:
        pass


class StochasticQuantClientEndpoint(QuantClientEndpoint):
    def __init__(self, **kwargs: Any) -> None:
        quant, dequant = stochastic_quantization(quantization_level=255)
        super().__init__(quant=quant, dequant=dequant, **kwargs)


class StochasticQuantServerEndpoint(QuantServerEndpoint):
    def __init__(self, **kwargs: Any) -> None:
        quant, dequant = stochastic_quantization(quantization_level=255)
        super().__init__(quant=quant, dequant=dequant, **kwargs)


class NNADQClientEndpoint(QuantClientEndpoint):
    def __init__(self, weight: float, **kwargs: Any) -> None:
        get_logger().info("use weight %s", weight)
        quant, dequant = NNADQ(weight=weight)
        super().__init__(quant=quant, dequant=dequant, **kwargs)

    def _after_quant(self, data: Any) -> None:
        NeuralNetworkAdaptiveDeterministicQuant.check_compression_ratio(quantized_data=data, prefix="worker")


class NNADQServerEndpoint(QuantServerEndpoint):
    def __init__(self, weight: float | None = None, **kwargs: Any) -> None:
        if weight is None:
            quant = None
            dequant: Callable = NeuralNetworkAdaptiveDeterministicDequant()
        else:
            quant, dequant = NNADQ(weight=weight)
        super().__init__(quant=quant, dequant=dequant, **kwargs)
        self.__quant_callback: Callable | None = None

    def set_quant_callback(self, quant_callback: Callable) -> None:
        self.__quant_callback = quant_callback

    def _after_quant(self, data: Any) -> None:
        NeuralNetworkAdaptiveDeterministicQuant.check_compression_ratio(data, prefix="broadcast")
        if self.__quant_callback is not None:
            self.__quant_callback(data=data)
