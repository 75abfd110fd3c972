"""Batched evaluation of coalition models for the Shapley engines.

The reference evaluates one coalition at a time: CPU weighted average of the members' models,
host->device load into the tester, a full eager test inference, offload
(simulation_lib/method/shapley_value/shapley_value_algorithm.py:67-76, server/server.py:40-55).
Here the N uploaded models are ``[P]`` rows in HBM and a whole list of coalitions is evaluated
together:

1. coalition models ``W_s = sum_i r_i^s W_i`` are NEVER materialised: every layer is one launch of
   the coalition GEMM (``ops.gemm.subset_linear``, csrc/subset_gemm.cu) whose B-operand producer
   warps read the members' fp32 layer slices straight from the uploaded arenas, form the weighted
   sum in registers and write the bf16 tile into the swizzled shared-memory slot the
   ``tcgen05.mma`` reads; the coalition bias is built in the epilogue the same way.  The only
   per-chunk host input is the ``[S, N]`` coefficient matrix;
2. the test set runs through all S coalitions at once: conv1 is ONE GEMM whose N dimension stacks
   the first-layer filters of all coalitions (the image patches are shared), conv2 and the three
   fully connected layers are one batched launch each with batch = coalition; patch extraction
   (unfold) and pooling stay in PyTorch.  (``subset_eval_fused: false`` selects the older path
   that first writes the S models into an ``[S, P]`` buffer with the reduce kernel and then runs
   ``linear_bf16``);
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

STATS = {"batched_calls": 0, "coalitions": 0, "fused_layer_launches": 0, "calls": 0, "seconds": 0.0}


_PROFILE = __import__("os").environ.get("DLS_SUBSET_PROFILE", "0") == "1"
STAGE_MS: dict = {}


class _StageTimer:
    """CUDA-event stage timing of the fused forward (DLS_SUBSET_PROFILE=1; synchronises, so profiling only)."""

    def __init__(self) -> None:
        self.events = [("start", self._event())]

    @staticmethod
    def _event() -> torch.cuda.Event:
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def __call__(self, name: str) -> None:
        self.events.append((name, self._event()))

    def flush(self) -> None:
        torch.cuda.synchronize()
        for (_, a), (name, b) in zip(self.events, self.events[1:]):
            STAGE_MS[name] = STAGE_MS.get(name, 0.0) + a.elapsed_time(b)


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
        # the coalition GEMM keeps the workers' arena pointers in shared memory: at most 64 of them
        self.fused = bool(server.config.algorithm_kwargs.get("subset_eval_fused", True)) and len(flats) <= 64
        # the fused path walks the test set in large slices (GEMM M dimension = samples x positions)
        self.eval_batch = int(server.config.algorithm_kwargs.get("subset_eval_batch", 1024))
        self._patch_cache: dict = {}
        self.worker_ids = sorted(flats)
        self.rows = [flats[w].flat for w in self.worker_ids]
        self.table = runtime.pointer_table(self.rows, self.device)
        self.index = {name: i for i, name in enumerate(layout.names)}

    def _coefficients(self, coalitions: list[tuple]) -> torch.Tensor:
        coefs = torch.zeros(len(coalitions), len(self.worker_ids), dtype=torch.float32)
        column = {w: j for j, w in enumerate(self.worker_ids)}
        for s, members in enumerate(coalitions):
            total = sum(self.sizes[w] for w in members)
            for w in members:
                coefs[s, column[w]] = self.sizes[w] / total
        return coefs.pin_memory().to(self.device, non_blocking=True)

    def _layer(self, x: torch.Tensor, coefs: torch.Tensor, name: str, stacked: bool, relu: bool,
               out_dtype: torch.dtype) -> torch.Tensor:
        from . import gemm

        wi, bi = self.index[f"{name}.weight"], self.index[f"{name}.bias"]
        shape = self.layout.shapes[wi]
        n_layer, k_w = int(shape[0]), int(self.layout.numels[wi]) // int(shape[0])
        STATS["fused_layer_launches"] += 1
        return gemm.subset_linear(x, self.rows, self.table, coefs, self.layout.offsets[wi], self.layout.offsets[bi],
                                  n_layer, k_w, stacked=stacked, relu=relu, out_dtype=out_dtype)

    def _forward_fused(self, x: torch.Tensor, coefs: torch.Tensor, cache_key: Any = None) -> torch.Tensor:
        """Logits [S, B, classes] of all coalitions; no coalition model exists in memory."""
        from . import misc

        F = torch.nn.functional
        mark = _StageTimer() if _PROFILE else None
        S, B = coefs.shape[0], x.shape[0]
        w1, w2 = self.index["conv1.weight"], self.index["conv2.weight"]
        k1, k2 = int(self.layout.shapes[w1][-1]), int(self.layout.shapes[w2][-1])
        c1, c2_in, c2 = int(self.layout.shapes[w1][0]), int(self.layout.shapes[w2][1]), int(self.layout.shapes[w2][0])
        hw = x.shape[-1] + 2 * self.pad - k1 + 1
        cols = self._patch_cache.get(cache_key) if cache_key is not None else None
        if cols is None:  # first-layer patches do not depend on the coalition: built once per test slice
            # own im2col (zero padding handled in the kernel), bf16, K padded to 8, the arena's filter order
            nhwc = x.to(torch.bfloat16).permute(0, 2, 3, 1).contiguous()
            cols = misc.im2col_grouped(nhwc, 1, k1, bool(self.layout.formats[w1]), pad=self.pad)[0]
            if cache_key is not None:
                self._patch_cache[cache_key] = cols
        if mark: mark("patches1")
        z = self._layer(cols, coefs, "conv1", True, True, torch.bfloat16)                        # [B*hw*hw, S*c1]
        if mark: mark("conv1_gemm")
        h = F.max_pool2d(z.view(B, hw, hw, S * c1).permute(0, 3, 1, 2), 2)                       # [B, S*c1, ., .]
        if mark: mark("pool1")
        hw2 = h.shape[-1] - k2 + 1
        # second-layer patches of all coalitions in one pass (csrc/misc.cu): [S, B*L, K], filter order of the arena
        cols = misc.im2col_grouped(h.permute(0, 2, 3, 1).contiguous(), S, k2, bool(self.layout.formats[w2]))
        if mark: mark("im2col2")
        z = self._layer(cols, coefs, "conv2", False, True, torch.bfloat16)                      # [S, B*L, c2]
        if mark: mark("conv2_gemm")
        h = z.view(S, B, hw2, hw2, c2).permute(0, 1, 4, 2, 3).reshape(S * B, c2, hw2, hw2)
        feat = F.max_pool2d(h, 2).reshape(S, B, -1)                                             # (c, h, w) order
        if mark: mark("pool2")
        z = self._layer(feat, coefs, "fc1", False, True, torch.bfloat16)
        z = self._layer(z, coefs, "fc2", False, True, torch.bfloat16)
        out = self._layer(z, coefs, "fc3", False, False, torch.float32)
        if mark:
            mark("fc")
            mark.flush()
        return out

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
        import time

        start = time.perf_counter()
        try:
            return self._evaluate(coalitions)
        finally:  # every chunk ends with a device->host read, so host time covers the device work
            STATS["calls"] += 1
            STATS["seconds"] += time.perf_counter() - start

    def _evaluate(self, coalitions: list[tuple]) -> list[float]:
        from . import gemm

        results: list[float] = []
        tester = self.server.tester
        tester.set_device(self.server._get_device(), stream=self.server._server_stream())
        phase = tester.phase
        F = torch.nn.functional
        for start in range(0, len(coalitions), self.max_coalitions):
            chunk = coalitions[start: start + self.max_coalitions]
            S = len(chunk)
            acc = torch.zeros(S, 3, dtype=torch.float64, device=self.device)
            if self.fused:
                coefs = self._coefficients(chunk)
                for key, x, y in self._test_slices(tester, phase):
                    logits = self._forward_fused(x, coefs, key)
                    self._accumulate(acc, logits, y)
                self._finish(tester, acc, results)
                continue
            models = self._coalition_models(chunk)
            p = {name: torch.stack([self.layout.view_of(models[s], i) for s in range(S)])
                 for i, name in enumerate(self.layout.names)}
            w1 = p["conv1.weight"].reshape(S * 6, *p["conv1.weight"].shape[2:])
            b1 = p["conv1.bias"].reshape(S * 6)
            w2 = p["conv2.weight"].reshape(S * 16, *p["conv2.weight"].shape[2:])
            b2 = p["conv2.bias"].reshape(S * 16)
            fw = [p[f"fc{j}.weight"].to(torch.bfloat16).contiguous() for j in (1, 2, 3)]
            fb = [p[f"fc{j}.bias"].contiguous() for j in (1, 2, 3)]
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
                self._accumulate(acc, logits, y)
            self._finish(tester, acc, results)
        return results

    def _test_slices(self, tester: Any, phase: Any) -> Any:
        """The tester's evaluation set in slices of ``subset_eval_batch`` samples (device resident;
        first-layer patches are cached per slice for the lifetime of this evaluator = one round)."""
        from ..engine.dataloader import SampleLoader

        loader = tester._loader(phase)
        if not isinstance(loader, SampleLoader) or loader.shuffle:
            for batch in loader:
                yield None, batch.inputs[0], batch.targets
            return
        cached = self._patch_cache.get("slices")
        if cached is None:
            x, y, order = loader.epoch_tensors()
            xs, ys = x.index_select(0, order), y.index_select(0, order)
            cached = [(("patches", i), xs[i: i + self.eval_batch], ys[i: i + self.eval_batch])
                      for i in range(0, ys.shape[0], self.eval_batch)]
            self._patch_cache["slices"] = cached
        yield from cached

    @staticmethod
    def _accumulate(acc: torch.Tensor, logits: torch.Tensor, y: torch.Tensor) -> None:
        S, B = logits.shape[0], logits.shape[1]
        loss = torch.nn.functional.cross_entropy(logits.reshape(S * B, -1), y.repeat(S),
                                                 reduction="none").view(S, B).sum(1)
        correct = (logits.argmax(-1) == y.unsqueeze(0)).sum(1)
        acc[:, 0] += loss.double()
        acc[:, 1] += correct.double()
        acc[:, 2] += B

    def _finish(self, tester: Any, acc: torch.Tensor, results: list[float]) -> None:
        reducer = getattr(tester, "_reducer", None)
        if reducer is not None:
            reducer.all_reduce_sum(acc)
        acc = acc.cpu()
        for s in range(acc.shape[0]):
            n = max(float(acc[s, 2]), 1.0)
            metric = {"loss": float(acc[s, 0]) / n, "accuracy": float(acc[s, 1]) / n}
            results.append(metric[self.metric_type])
        STATS["batched_calls"] += 1
        STATS["coalitions"] += acc.shape[0]
