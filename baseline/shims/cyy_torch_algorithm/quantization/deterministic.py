import math

import torch
from cyy_naive_lib.log import get_logger


class NeuralNetworkAdaptiveDeterministicQuant:
    def __init__(self, weight: float) -> None:
        self.weight = weight

    def __call__(self, data: dict) -> dict:
        out = {}
        for k, x in data.items():
            mn, rng, sd = x.min(), x.max() - x.min(), x.float().std(unbiased=False)
            b = 1
            if float(rng) > 0:
                c = math.log(2.0) * float(rng) / (math.sqrt(12.0) * max(float(sd), 1e-30) * self.weight)
                b = int(min(16, max(1, round(math.log2(max(c, 1e-30))))))
            step = float(rng) / (2 ** b - 1) if float(rng) > 0 else 1.0
            code = ((x - mn) / step).round().clamp(0, 2 ** b - 1)
            out[k] = {"code": code.to(torch.int32 if b > 8 else torch.uint8), "min": mn, "step": step, "bits": b}
        return out

    @classmethod
    def check_compression_ratio(cls, quantized_data, prefix: str) -> float:
        payload = getattr(quantized_data, "parameter", quantized_data)
        bits = sum(v["code"].numel() * v["bits"] for v in payload.values() if isinstance(v, dict))
        total = sum(v["code"].numel() * 32 for v in payload.values() if isinstance(v, dict))
        ratio = bits / max(total, 1)
        get_logger().info("%s NNABQ compression ratio is %s", prefix, ratio)
        return ratio


class NeuralNetworkAdaptiveDeterministicDequant:
    def __call__(self, data):
        if not isinstance(data, dict):
            return data
        return {k: (v["code"].float() * v["step"] + v["min"]) if isinstance(v, dict) and "code" in v else v
                for k, v in data.items()}


def NNADQ(weight: float):
    return NeuralNetworkAdaptiveDeterministicQuant(weight), NeuralNetworkAdaptiveDeterministicDequant()
