"""Batched evaluation of coalition models for the Shapley engines.

The reference evaluates one coalition at a time: CPU weighted average of the members' models,
host->device load into the tester, a full eager test inference, offload
(simulation_lib/method/shapley_value/shapley_value_algorithm.py:67-76, server/server.py:40-55).
Here the N uploaded models are ``[P]`` rows in HBM and a whole list of coalitions is evaluated
together:

1. coalition models ``W_s = sum_i r_i^s W_i`` are built by the fused reduce kernel straight
   into one ``[S, P]`` buffer (no per-coalition tensors, no host round trip);
2. the test set runs through all S models at once, every matrix product on the hand-written
   tcgen05 GEMM (``ops.gemm.linear_bf16``, bias and ReLU fused in its epilogue): conv1 is ONE GEMM
   whose B operand stacks the first-layer filters of all coalitions (the image patches are shared,
   so the coalition models are never materialised as separate convolutions), conv2 and the three
   fully connected layers are one batched launch each with batch = coalition; patch extraction
   (unfold) and pooling stay in PyTorch;
3. per-coalition accuracy / loss are reduced on the device and read back once.

The GTG engine's truncation rules are applied afterwards to decide which utilities are
*used*, so the Shapley values equal the sequential evaluation's up to bf16 rounding of the
fully connected layers (argmax metric: a handful of borderline predictions out of thousands).
Supported tester models: ``LeNet5`` (the reference's GTG/multi-round/hierarchical configs,
conf/gtg_sv/mnist.yaml:3); anything else returns ``None`` and is evaluated one by one.
"""
from typing import Any

import torch

from ..engine.model_util import as_flat
from . import runtime
from .reduce import reduce_rows_f32

STATS = {"batched_calls": 0, "coalitions": 0}


def make_batch_evaluator(server: Any, all_worker_data: dict, metric_type: str) -> Any:
    if not runtime.available() or not server.config.algorithm_kwargs.get("batched_subset_eval", True):
        return None
    from ..message import ParameterMessage
    from ..models.vision import LeNet5

    tester = server.tester
    if not isinstance(tester.model, LeNet5):
        return None
    flats = {}
    for worker_id, msg in all_worker_data.items():
        flat = as_flat(msg.parameter) if isinstance(msg, ParameterMessage) else None
        if flat is None or not flat.flat.is_cuda:
            return None
        flats[worker_id] = flat
    layout = next(iter(flats.values())).layout
    if any(f.layout != layout for f in flats.values()):
        return None
    sizes = {w: float(m.dataset_size) for w, m in all_worker_data.items()}
    evaluator = _LeNet5BatchEvaluator(server, flats, sizes, layout, metric_type)
    return evaluator


class _LeNet5BatchEvaluator:
    def __init__(self, server: Any, flats: dict, sizes: dict, layout: Any, metric_type: str) -> None:
        self.server = server
        self.flats = flats
        self.sizes = sizes
        self.layout = layout
        self.metric_type = metric_type
        self.device = next(iter(flats.values())).flat.device
        self.max_coalitions = int(server.config.algorithm_kwargs.get("subset_eval_chunk", 32))
        self.pad = int(server.tester.model.conv1.padding[0])

    def _coalition_models(self, coalitions: list[tuple]) -> torch.Tensor:
        # rows are padded to 256 bytes: the reduce kernel uses 128-bit accesses
        stride = (self.layout.total + 63) // 64 * 64
        out = torch.empty(len(coalitions), stride, dtype=torch.float32, device=self.device)[:, : self.layout.total]
        for s, members in enumerate(coalitions):
            total = sum(self.sizes[w] for w in members)
            rows = [self.flats[w].flat for w in members]
            weights = [self.sizes[w] / total for w in members]
            reduce_rows_f32(rows, weights, None, 0.0, 1.0, out=out[s], check_nan=False)
        return out

    @torch.no_grad()
    def __call__(self, coalitions: list[tuple]) -> list[float]:
        from . import gemm

        results: list[float] = []
        tester = self.server.tester
        tester.set_device(self.server._get_device(), stream=self.server._server_stream())
        phase = tester.phase
        F = torch.nn.functional
        for start in range(0, len(coalitions), self.max_coalitions):
            chunk = coalitions[start: start + self.max_coalitions]
            S = len(chunk)
            models = self._coalition_models(chunk)
            p = {name: torch.stack([self.layout.view_of(models[s], i) for s in range(S)])
                 for i, name in enumerate(self.layout.names)}
            w1 = p["conv1.weight"].reshape(S * 6, *p["conv1.weight"].shape[2:])
            b1 = p["conv1.bias"].reshape(S * 6)
            w2 = p["conv2.weight"].reshape(S * 16, *p["conv2.weight"].shape[2:])
            b2 = p["conv2.bias"].reshape(S * 16)
            fw = [p[f"fc{j}.weight"].to(torch.bfloat16).contiguous() for j in (1, 2, 3)]
            fb = [p[f"fc{j}.bias"].contiguous() for j in (1, 2, 3)]
            acc = torch.zeros(S, 3, dtype=torch.float64, device=self.device)
            conv_gemm = bool(self.server.config.algorithm_kwargs.get("subset_eval_conv_gemm", True))
            if conv_gemm:
                # convolution weights as GEMM operands: conv1 of ALL coalitions is one B matrix
                # [S*6, C*k*k] (the image patches are shared), conv2 is one [16, 6*k*k] matrix per coalition
                k1, k2 = w1.shape[-1], w2.shape[-1]
                g1 = w1.reshape(S * 6, -1).to(torch.bfloat16).contiguous()
                g2 = p["conv2.weight"].reshape(S, 16, -1).to(torch.bfloat16).contiguous()
            for batch in tester._loader(phase):
                x, y = batch.inputs[0], batch.targets
                B = x.shape[0]
                if conv_gemm:
                    hw = x.shape[-1] + 2 * self.pad - k1 + 1
                    cols = F.unfold(x, k1, padding=self.pad).transpose(1, 2).reshape(B * hw * hw, -1)
                    z = gemm.linear_bf16(cols.to(torch.bfloat16), g1, b1, relu=True, out_dtype=torch.bfloat16)
                    h = z.view(B, hw, hw, S * 6).permute(0, 3, 1, 2)            # channels-last [B, S*6, 28, 28]
                    h = F.max_pool2d(h, 2)                                        # [B, S*6, 14, 14]
                    hw2 = h.shape[-1] - k2 + 1
                    cols = F.unfold(h.float(), k2)                                 # [B, S*6*k*k, hw2*hw2]
                    cols = cols.view(B, S, 6 * k2 * k2, hw2 * hw2).permute(1, 0, 3, 2).reshape(S, B * hw2 * hw2, -1)
                    z = gemm.linear_bf16(cols.to(torch.bfloat16), g2, p["conv2.bias"].contiguous(), relu=True,
                                         out_dtype=torch.bfloat16)                # [S, B*100, 16]
                    h = z.view(S, B, hw2, hw2, 16).permute(0, 1, 4, 2, 3).reshape(S * B, 16, hw2, hw2)
                    h = F.max_pool2d(h.float(), 2)                                # [S*B, 16, 5, 5]
                    feat = h.reshape(S, B, -1).to(torch.bfloat16).contiguous()    # [S, B, 400] in (c, h, w) order
                else:
                    h = F.max_pool2d(F.relu(F.conv2d(x, w1, b1, padding=self.pad)), 2)      # [B, S*6, 14, 14]
                    h = F.max_pool2d(F.relu(F.conv2d(h, w2, b2, groups=S)), 2)                 # [B, S*16, 5, 5]
                    feat = h.reshape(B, S, -1).permute(1, 0, 2).to(torch.bfloat16).contiguous()  # [S, B, 400]
                z = gemm.linear_bf16(feat, fw[0], fb[0], relu=True, out_dtype=torch.bfloat16)
                z = gemm.linear_bf16(z, fw[1], fb[1], relu=True, out_dtype=torch.bfloat16)
                logits = gemm.linear_bf16(z, fw[2], fb[2], relu=False, out_dtype=torch.float32)  # [S, B, classes]
                loss = F.cross_entropy(logits.reshape(S * B, -1), y.repeat(S), reduction="none").view(S, B).sum(1)
                correct = (logits.argmax(-1) == y.unsqueeze(0)).sum(1)
                acc[:, 0] += loss.double()
                acc[:, 1] += correct.double()
                acc[:, 2] += B
            reducer = getattr(tester, "_reducer", None)
            if reducer is not None:
                reducer.all_reduce_sum(acc)
            acc = acc.cpu()
            for s in range(S):
                n = max(float(acc[s, 2]), 1.0)
                metric = {"loss": float(acc[s, 0]) / n, "accuracy": float(acc[s, 1]) / n}
                results.append(metric[self.metric_type])
            STATS["batched_calls"] += 1
            STATS["coalitions"] += S
        return results
