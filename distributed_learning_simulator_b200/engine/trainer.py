"""Trainer: local SGD with hook points.

Own version of ``cyy_torch_toolbox.Trainer`` with the contract the reference uses
(/root/reference/simulation_lib/worker/worker.py:67-89,
worker/aggregation_worker.py:25-130, worker/gradient_worker.py:28-131,
util/model.py:6-23).  Two execution paths share the same hooks:

* eager path (CPU, or any run with per-batch hooks such as ``OPTIMIZER_STEP``): a plain
  PyTorch loop;
* B200 path (``engine/graph_step.py``): parameters and gradients live in a flat fp32
  arena, the optimizer is a hand-written fused SGD kernel that reads the learning rate
  from a device-side schedule table, and forward+backward+update of one batch is
  captured once in a CUDA graph and replayed per batch on the worker's own stream.
"""
import time
import warnings
from typing import Any

import torch
from torch import nn

from .dataset import DatasetCollection
from .device import capture_guard
from .executor import HookConfig, ModelExecutor
from .hyper_parameter import HyperParameter
from .log import get_logger
from .ml_type import ExecutorHookPoint, MachineLearningPhase, StopExecutingException
from .tensor import TensorDict


class Trainer(ModelExecutor):
    def __init__(self, model: nn.Module, dataset_collection: DatasetCollection, hyper_parameter: HyperParameter,
                 hook_config: HookConfig | None = None) -> None:
        super().__init__(model, dataset_collection, MachineLearningPhase.Training, hyper_parameter, hook_config)
        self._optimizer: torch.optim.Optimizer | None = None
        self._lr_scheduler: Any = None
        self._graph_step: Any = None
        self.limited_resource = False
        self.trained_sample_count = 0
        self.last_train_seconds = 0.0

    # ---- optimizer ---------------------------------------------------------------
    def get_optimizer(self) -> torch.optim.Optimizer:
        if self._optimizer is None:
            self._optimizer = self._hyper_parameter.get_optimizer(list(self._model.parameters()))
        return self._optimizer

    def remove_optimizer(self) -> None:
        self._optimizer = None
        self._lr_scheduler = None
        if self._graph_step is not None:
            self._graph_step.reset_optimizer_state()

    def _get_lr_scheduler(self, total_steps: int) -> Any:
        if self._lr_scheduler is None:
            self._lr_scheduler = self._hyper_parameter.get_lr_scheduler(self.get_optimizer(), total_steps)
        return self._lr_scheduler

    def load_parameter_dict(self, parameter_dict: TensorDict) -> None:
        """Load parameters and start from a fresh optimizer/scheduler
        (reference: util/model.py:21-22)."""
        self._model_util.load_parameter_dict(parameter_dict)
        self.remove_optimizer()
        self.exec_hooks(ExecutorHookPoint.MODEL_LOADED, executor=self)

    def load_parameter_list(self, flat: torch.Tensor, reuse_learning_rate: bool = False) -> None:
        self._model_util.load_parameter_list(flat)
        if not reuse_learning_rate:
            self.remove_optimizer()
        elif self._optimizer is not None:
            self._optimizer.state.clear()
        if reuse_learning_rate and self._graph_step is not None:
            self._graph_step.reset_optimizer_state(keep_schedule=True)
        self.exec_hooks(ExecutorHookPoint.MODEL_LOADED, executor=self)

    def _on_device_change(self) -> None:
        self._optimizer = None
        self._lr_scheduler = None
        self._graph_step = None

    # ---- training ------------------------------------------------------------------
    def _needs_eager(self) -> bool:
        if self._device.type != "cuda" or not self.hook_config.use_cuda_graph:
            return True
        if self._dc.is_graph:
            return True
        per_batch = (ExecutorHookPoint.OPTIMIZER_STEP, ExecutorHookPoint.BEFORE_BATCH,
                     ExecutorHookPoint.AFTER_FORWARD, ExecutorHookPoint.AFTER_BATCH)
        return any(self.has_hook(p) for p in per_batch)

    def train(self, **kwargs: Any) -> None:
        start = time.monotonic()
        self._model.train()
        self._metric.clear()
        try:
            with self._stream_context():
                self.exec_hooks(ExecutorHookPoint.BEFORE_EXECUTE, executor=self)
                if self._needs_eager():
                    self._train_eager()
                else:
                    self._train_graph()
                self.exec_hooks(ExecutorHookPoint.AFTER_EXECUTE, executor=self, epoch=self._hyper_parameter.epoch)
        except StopExecutingException:
            get_logger().debug("training stopped by a hook")
        self.last_train_seconds = time.monotonic() - start

    def _epoch_end(self, epoch: int, epoch_start: float) -> None:
        m = self._metric.end_epoch(epoch)
        self.trained_sample_count += int(m["sample_count"])
        if self.hook_config.log_performance_metric:
            get_logger().info("%sepoch: %d, train in %.3f s, learning rate: %e, loss: %e, accuracy: %.2f%%",
                              self._visualizer_prefix, epoch, time.monotonic() - epoch_start, self.current_lr(),
                              float(m["loss"]), float(m["accuracy"]) * 100)
        if self._dc.has_dataset(MachineLearningPhase.Validation) and self.has_hook(ExecutorHookPoint.AFTER_VALIDATION):
            vm = self._evaluate(MachineLearningPhase.Validation)
            if self.hook_config.log_performance_metric:
                get_logger().info("%sepoch: %d, validation loss: %e, accuracy: %.2f%%", self._visualizer_prefix,
                                  epoch, vm["loss"], vm["accuracy"] * 100)
            self.exec_hooks(ExecutorHookPoint.AFTER_VALIDATION, executor=self, epoch=epoch, validation_metric=vm)
        self.exec_hooks(ExecutorHookPoint.AFTER_EPOCH, executor=self, epoch=epoch)

    def current_lr(self) -> float:
        if self._graph_step is not None and self._graph_step.active:
            return self._graph_step.current_lr()
        if self._optimizer is not None:
            return float(self._optimizer.param_groups[0]["lr"])
        return self._hyper_parameter.learning_rate

    def _train_eager(self) -> None:
        self._model_util.disable_shadow()  # torch.optim updates fp32 parameters
        optimizer = self.get_optimizer()
        loader = self._loader(MachineLearningPhase.Training)
        scheduler = self._get_lr_scheduler(self._hyper_parameter.epoch * len(loader))
        if self._model_util.flat_gradient is not None:
            self._model_util.rebind_flat_gradients()
        for epoch in range(1, self._hyper_parameter.epoch + 1):
            epoch_start = time.monotonic()
            self.exec_hooks(ExecutorHookPoint.BEFORE_EPOCH, executor=self, epoch=epoch)
            self._metric.start_epoch(self._device)
            for batch_index, batch in enumerate(loader):
                self.exec_hooks(ExecutorHookPoint.BEFORE_BATCH, executor=self, epoch=epoch, batch_index=batch_index)
                # library-backed eager work holds the capture guard shared, one batch at a time and never
                # across a hook that may wait for another role (a pending capture blocks new sharers)
                with capture_guard.shared():
                    optimizer.zero_grad(set_to_none=self._model_util.flat_gradient is None)
                    output, loss = self._forward(batch)
                self.exec_hooks(ExecutorHookPoint.AFTER_FORWARD, executor=self, epoch=epoch, batch_index=batch_index)
                with capture_guard.shared():
                    loss.backward()
                if self.has_hook(ExecutorHookPoint.OPTIMIZER_STEP):
                    # the hook owns the update (sign_SGD applies the server's majority vote), so the
                    # scheduler legitimately ticks without torch having seen optimizer.step()
                    self.exec_hooks(ExecutorHookPoint.OPTIMIZER_STEP, executor=self)
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore", UserWarning)
                        scheduler.step()
                else:
                    with capture_guard.shared():
                        optimizer.step()
                    scheduler.step()
                self._metric.add_batch(loss, output, batch.targets)
                self.exec_hooks(ExecutorHookPoint.AFTER_BATCH, executor=self, epoch=epoch, batch_index=batch_index,
                                batch_size=batch.batch_size)
            self._epoch_end(epoch, epoch_start)

    def _train_graph(self) -> None:
        from .graph_step import GraphTrainStep

        if self._graph_step is None:
            self._graph_step = GraphTrainStep(self)
        self._graph_step.run()
