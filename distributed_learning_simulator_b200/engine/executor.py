"""Shared base of Trainer and Inferencer (the reference's ``cyy_torch_toolbox``
"executor" concept: a model bound to a dataset collection, a phase, a device, hooks and
a performance metric; contract reconstructed from the call sites listed in SURVEY.md 2.3).
"""
import os
from dataclasses import dataclass
from typing import Any

import torch
from torch import nn

from .dataloader import Batch, create_loader
from .dataset import DatasetCollection
from .device import capture_guard, capture_stream
from .hook import HookCollection
from .hyper_parameter import HyperParameter
from .metric import PerformanceMetric
from .ml_type import MachineLearningPhase
from .model_util import ModelUtil
from ..ops import conv as conv_ops


@dataclass
class HookConfig:
    use_amp: bool = False
    summarize_executor: bool = True
    log_performance_metric: bool = True
    debug: bool = False
    use_cuda_graph: bool = True   # B200 path: capture the train step
    use_fused_kernels: bool = True  # B200 path: flat-arena fused SGD & friends


class ModelExecutor(HookCollection):
    def __init__(self, model: nn.Module, dataset_collection: DatasetCollection, phase: MachineLearningPhase,
                 hyper_parameter: HyperParameter, hook_config: HookConfig | None = None) -> None:
        super().__init__()
        self._model = model
        conv_ops.install(model)  # nn.Conv2d -> this framework's kernels on CUDA (no-op elsewhere)
        self._model_util = ModelUtil(model)
        self._dc = dataset_collection
        self._phase = phase
        self._hyper_parameter = hyper_parameter
        self.hook_config = hook_config or HookConfig()
        self._device = torch.device("cpu")
        self._stream: torch.cuda.Stream | None = None
        self._metric = PerformanceMetric()
        self._visualizer_prefix = ""
        self._dataloader_kwargs: dict = {}
        self._generator: torch.Generator | None = None
        self._seed = 0
        self.loss_fun = nn.CrossEntropyLoss()

    # ---- properties the reference reads ------------------------------------------
    @property
    def model(self) -> nn.Module:
        return self._model

    @property
    def model_util(self) -> ModelUtil:
        return self._model_util

    @property
    def dataset_collection(self) -> DatasetCollection:
        return self._dc

    @property
    def phase(self) -> MachineLearningPhase:
        return self._phase

    @property
    def hyper_parameter(self) -> HyperParameter:
        return self._hyper_parameter

    @property
    def performance_metric(self) -> PerformanceMetric:
        return self._metric

    @property
    def device(self) -> torch.device:
        return self._device

    @property
    def dataset_size(self) -> int:
        return self._dc.get_dataset_size(self._phase)

    @property
    def stream(self) -> torch.cuda.Stream | None:
        return self._stream

    def set_visualizer_prefix(self, prefix: str) -> None:
        self._visualizer_prefix = prefix

    def update_dataloader_kwargs(self, **kwargs: Any) -> None:
        self._dataloader_kwargs.update(kwargs)

    def set_seed(self, seed: int) -> None:
        self._seed = seed
        self._generator = None

    # ---- device ------------------------------------------------------------------
    def set_device(self, device: torch.device, stream: "torch.cuda.Stream | None" = None) -> None:
        if device == self._device and (device.type != "cuda" or self._stream is not None):
            self._model_util.to_device(device)
            self._model_util.flatten_parameters(with_grad=False)
            if stream is not None:
                self._stream = stream
            return
        self._device = device
        self._model_util.to_device(device)
        self._model_util.flatten_parameters(with_grad=False)
        self._generator = None
        if device.type == "cuda":
            # one stream per executor: co-located workers overlap on the GPU
            self._stream = stream if stream is not None else torch.cuda.Stream(device=device)
        else:
            self._stream = None
        self._on_device_change()

    def stream_context(self) -> Any:
        return self._stream_context()

    def _on_device_change(self) -> None:
        pass

    def offload_from_device(self) -> None:
        """The reference moves the model to the CPU between rounds to fit many workers
        on a small GPU (worker/aggregation_worker.py:57,124-130).  With 180 GB of HBM3e
        the model stays resident; only an explicit ``limited_resource`` run offloads."""
        if getattr(self, "limited_resource", False) and self._device.type == "cuda":
            self.wait_stream()
            self._model_util.to_device(torch.device("cpu"))

    def wait_stream(self) -> None:
        if self._stream is not None:
            self._stream.synchronize()

    def _stream_context(self) -> Any:
        if self._stream is not None:
            return torch.cuda.stream(self._stream)
        return _NullContext()

    def _rng(self) -> torch.Generator:
        if self._generator is None or self._generator.device != self._device:
            self._generator = torch.Generator(device=self._device)
            self._generator.manual_seed(self._seed)
        return self._generator

    def _autocast(self) -> Any:
        if self.hook_config.use_amp:
            return torch.autocast(device_type=self._device.type, dtype=torch.bfloat16)
        return _NullContext()

    # ---- forward -----------------------------------------------------------------
    def _forward(self, batch: Batch) -> tuple[torch.Tensor, torch.Tensor]:
        with self._autocast():
            output = self._model(*batch.inputs, **{k: v for k, v in batch.kwargs.items() if k != "output_index"})
            if "output_index" in batch.kwargs:
                output = output[batch.kwargs["output_index"]]
            elif batch.output_rows is not None:
                output = output[: batch.output_rows]
            loss = self.loss_fun(output.float(), batch.targets)
        return output, loss

    # ---- evaluation as CUDA-graph replays ----------------------------------------------------------
    def _graph_eval_batch(self, batch: Batch, metric: PerformanceMetric) -> bool:
        """Evaluates one image batch by replaying a captured forward + loss + metric-accumulation graph (one
        graph per batch shape; the eager path costs ~25 us of host time per kernel, ~3 ms per batch).  Returns
        False when the batch does not qualify -- the caller then takes the eager path."""
        if self._device.type != "cuda" or not self.hook_config.use_cuda_graph or not self.hook_config.use_fused_kernels:
            return False
        if os.environ.get("DLS_EVAL_GRAPH", "1") == "0" or len(batch.inputs) != 1 or batch.kwargs \
                or batch.output_rows is not None:
            return False
        x = batch.inputs[0]
        if not (isinstance(x, torch.Tensor) and x.dim() == 4 and x.is_floating_point() and x.is_cuda):
            return False
        from ..ops import runtime
        from ..ops import sgd as sgd_ops

        if not runtime.available():
            return False
        if self.__dict__.get("_eval_graph_broken", False):
            return False
        graphs = self.__dict__.setdefault("_eval_graphs", {})
        first = next(self._model.parameters(), None)
        key = (tuple(x.shape), x.dtype, tuple(x.stride()), bool(self.hook_config.use_amp),
               None if first is None else (first.data_ptr(), first.dtype))  # a re-homed model gets a fresh capture
        entry = graphs.get(key)
        if entry is None:
            static_x = torch.empty_strided(x.shape, x.stride(), dtype=x.dtype, device=x.device)
            static_y = torch.empty_like(batch.targets)
            acc = torch.zeros(3, dtype=torch.float64, device=x.device)

            stats_floats = sum(2 * m.out_channels for m in self._model.modules() if type(m) is nn.Conv2d)
            stats_buf = torch.zeros(max(stats_floats, 1), dtype=torch.float32, device=x.device)

            def run() -> None:
                stats_buf.zero_()  # BatchNorm statistics of the convolution epilogues: one fill per batch
                with self._autocast(), conv_ops.stats_arena(stats_buf):
                    output = self._model(static_x)
                    loss = self.loss_fun(output.float(), static_y)
                sgd_ops.metric_accumulate(output.float().contiguous(), static_y, loss.detach().reshape(1), acc)

            try:
                with capture_guard.exclusive():
                    static_x.copy_(x)
                    static_y.copy_(batch.targets)
                    side = torch.cuda.Stream(device=self._device)
                    side.wait_stream(torch.cuda.current_stream(self._device))
                    with torch.cuda.stream(side):
                        for _ in range(2):
                            run()
                    torch.cuda.current_stream(self._device).wait_stream(side)
                    torch.cuda.synchronize(self._device)
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=capture_stream(self._device), capture_error_mode="thread_local"):
                        run()
                    torch.cuda.synchronize(self._device)
            except RuntimeError as e:
                # a capture that another role's eager library call invalidated: this executor evaluates eagerly
                # from now on (correct, just ~3 ms of host time per batch slower)
                from .log import get_logger

                get_logger().warning("evaluation graph capture failed, falling back to eager evaluation: %s",
                                     str(e).splitlines()[0])
                self.__dict__["_eval_graph_broken"] = True
                torch.cuda.synchronize(self._device)
                return False
            # everything the captured kernels address must live as long as the graph (stats_buf included)
            entry = graphs[key] = (graph, static_x, static_y, acc, stats_buf)
        graph, static_x, static_y, acc, _ = entry
        static_x.copy_(x, non_blocking=True)
        static_y.copy_(batch.targets, non_blocking=True)
        acc.zero_()
        graph.replay()
        metric.accumulator.add_(acc)
        return True

    def _loader(self, phase: MachineLearningPhase) -> Any:
        batch_size = self._hyper_parameter.batch_size
        if phase != MachineLearningPhase.Training:
            batch_size = max(batch_size, int(self._hyper_parameter.extra_hyper_parameters.get(
                "inference_batch_size", 1024)))
        return create_loader(self._dc, phase, batch_size, self._device, model=self._model,
                             generator=self._rng(), **self._dataloader_kwargs)

    def _evaluate(self, phase: MachineLearningPhase, reducer: Any = None) -> dict[str, float]:
        """Loss/accuracy of the current model on ``phase`` (no grad, train-mode BN is
        irrelevant because running stats are disabled: batch statistics are always used).
        ``reducer`` (a distributed context) sums the counters over ranks first."""
        metric = PerformanceMetric()
        was_training = self._model.training
        self._model.eval()
        with self._stream_context():
            self._model_util.refresh_shadow()  # the fp32 arena may have been written in place since
        with torch.no_grad(), self._stream_context():
            metric.start_epoch(self._device)
            if self._dc.get_dataset_size(phase) > 0:
                for batch in self._loader(phase):
                    if self._graph_eval_batch(batch, metric):
                        continue
                    with capture_guard.shared():
                        output, loss = self._forward(batch)
                        metric.add_batch(loss, output, batch.targets)
            if reducer is not None:  # a collective with other ranks: never while holding the guard
                reducer.all_reduce_sum(metric.accumulator)
            metric.end_epoch(1)
        self._model.train(was_training)
        return metric.get_epoch_metrics(1)


class _NullContext:
    def __enter__(self) -> None:
        return None

    def __exit__(self, *args: Any) -> None:
        return None
