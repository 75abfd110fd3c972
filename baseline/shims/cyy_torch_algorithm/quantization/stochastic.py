import torch


def stochastic_quantization(quantization_level: int = 255):
    def quant_one(x: torch.Tensor) -> dict:
        s = x.abs().max()
        y = x.abs() / s.clamp(min=1e-38) * quantization_level
        lo = y.floor()
        q = lo + (torch.rand_like(y) < (y - lo)).float()
        return {"q": q.to(torch.uint8), "sign": (x < 0), "norm": s, "level": quantization_level}

    def quant(data):
        return {k: quant_one(v) for k, v in data.items()}

    def dequant(data):
        if not isinstance(data, dict):
            return data
        out = {}
        for k, v in data.items():
            if isinstance(v, dict) and "q" in v:
                mag = v["q"].float() * v["norm"] / v["level"]
                out[k] = torch.where(v["sign"], -mag, mag)
            else:
                out[k] = v
        return out

    return quant, dequant
