"""Plain eager PyTorch trainer / inferencer with the hook surface the reference uses
(SURVEY.md 2.3): data cached on the CPU, one host->device copy per batch, torch.optim.SGD +
CosineAnnealingLR stepped per batch, bf16 autocast when ``use_amp`` is set, metrics summed on
the device and read once per epoch."""
import math
import os
import time

import torch
from cyy_naive_lib.log import get_logger
from torch import nn

from .ml_type import ExecutorHookPoint, MachineLearningPhase, StopExecutingException
from .model.util import ModelUtil


class HookConfig:
    def __init__(self) -> None:
        self.use_amp = False
        self.summarize_executor = True
        self.log_performance_metric = True


class HyperParameter:
    def __init__(self) -> None:
        self.epoch = 1
        self.batch_size = 64
        self.learning_rate = 0.01
        self.weight_decay = 0.0
        self.momentum = 0.9
        self.optimizer_name = "SGD"
        self.learning_rate_scheduler_name = "CosineAnnealingLR"
        self.extra_hyper_parameters: dict = {}


class PerformanceMetric:
    def __init__(self) -> None:
        self._epochs: dict = {}

    def set(self, epoch: int, loss: float, accuracy: float) -> None:
        self._epochs[epoch] = {"loss": torch.tensor(loss), "accuracy": torch.tensor(accuracy)}

    def get_epoch_metrics(self, epoch: int) -> dict:
        return {k: float(v) for k, v in self._epochs[epoch].items()}

    def get_loss(self, epoch: int):
        return self._epochs[epoch]["loss"]

    def get_accuracy(self, epoch: int):
        return self._epochs[epoch]["accuracy"]


class Executor:
    def __init__(self, model: nn.Module, dataset_collection, phase, hyper_parameter: HyperParameter,
                 hook_config: HookConfig) -> None:
        self.model = model
        self.model_util = ModelUtil(model)
        self.dataset_collection = dataset_collection
        self.phase = phase
        self.hyper_parameter = hyper_parameter
        self.hook_config = hook_config
        self.performance_metric = PerformanceMetric()
        self.device = torch.device("cpu")
        self._hooks: dict = {}
        self._hook_objs: dict = {}
        self._disabled: set = set()
        self._prefix = ""
        self.loss_fun = nn.CrossEntropyLoss()

    # hooks -------------------------------------------------------------------------------
    def append_named_hook(self, hook_point, name: str, fun) -> None:
        self._hooks.setdefault(hook_point, []).append((name, fun))

    def remove_named_hook(self, name: str, hook_point=None) -> None:
        for point, entries in self._hooks.items():
            if hook_point is None or hook_point == point:
                entries[:] = [(n, f) for n, f in entries if n != name]

    def append_hook(self, hook, hook_name: str | None = None) -> None:
        hook_name = hook_name or type(hook).__name__
        self._hook_objs[hook_name] = hook
        for point, fn in hook.yield_hooks():
            self.append_named_hook(point, f"{hook_name}.{point.name}", fn)

    def remove_hook(self, hook_name: str) -> None:
        hook = self._hook_objs.pop(hook_name, None)
        if hook is not None:
            for point, _ in hook.yield_hooks():
                self.remove_named_hook(f"{hook_name}.{point.name}", point)

    def has_hook_obj(self, hook_name: str) -> bool:
        return hook_name in self._hook_objs

    def get_hook(self, hook_name: str):
        return self._hook_objs[hook_name]

    def disable_hook(self, hook_name: str) -> None:
        self._disabled.add(hook_name)

    def has_hook(self, hook_point) -> bool:
        return bool(self._hooks.get(hook_point))

    def exec_hooks(self, hook_point, **kwargs) -> None:
        for name, fun in list(self._hooks.get(hook_point, ())):
            if name.split(".")[0] not in self._disabled:
                fun(**kwargs)

    # device ----------------------------------------------------------------------------------
    def set_device(self, device) -> None:
        self.device = device
        self.model.to(device)

    def offload_from_device(self) -> None:
        self.model.to("cpu")
        if torch.cuda.is_available():
            torch.cuda.empty_cache()

    def wait_stream(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()

    def set_visualizer_prefix(self, prefix: str) -> None:
        self._prefix = prefix

    def update_dataloader_kwargs(self, **kwargs) -> None:
        pass

    @property
    def dataset_size(self) -> int:
        return self.dataset_collection.get_dataset_size(self.phase)

    # data -------------------------------------------------------------------------------------
    def _batches(self, phase, batch_size: int, shuffle: bool):
        p = self.dataset_collection.get_phase(phase)
        idx = p.indices
        if shuffle:
            idx = idx[torch.randperm(idx.numel())]
        for start in range(0, idx.numel(), batch_size):
            sel = idx[start: start + batch_size]
            yield (p.x[sel].to(self.device, non_blocking=True), p.y[sel].to(self.device, non_blocking=True))

    def _autocast(self):
        return torch.autocast(device_type=self.device.type, dtype=torch.bfloat16, enabled=bool(self.hook_config.use_amp))

    def _evaluate(self, phase) -> dict:
        self.model.to(self.device)
        acc = torch.zeros(3, dtype=torch.float64, device=self.device)
        with torch.no_grad():
            for x, y in self._batches(phase, max(self.hyper_parameter.batch_size, 1024), shuffle=False):
                with self._autocast():
                    out = self.model(x)
                    loss = self.loss_fun(out.float(), y)
                acc[0] += loss.double() * y.shape[0]
                acc[1] += (out.argmax(-1) == y).sum()
                acc[2] += y.shape[0]
        acc = acc.cpu()
        n = max(float(acc[2]), 1.0)
        return {"loss": float(acc[0]) / n, "accuracy": float(acc[1]) / n}


class Trainer(Executor):
    def __init__(self, model, dataset_collection, hyper_parameter, hook_config) -> None:
        super().__init__(model, dataset_collection, MachineLearningPhase.Training, hyper_parameter, hook_config)
        self._optimizer = None

    def get_optimizer(self):
        if self._optimizer is None:
            hp = self.hyper_parameter
            self._optimizer = torch.optim.SGD(self.model.parameters(), lr=hp.learning_rate, momentum=hp.momentum,
                                              weight_decay=hp.weight_decay)
        return self._optimizer

    def remove_optimizer(self) -> None:
        self._optimizer = None

    def load_parameter_dict(self, parameter_dict: dict) -> None:
        self.model_util.load_parameter_dict(parameter_dict)
        self.remove_optimizer()

    def train(self, **kwargs) -> None:
        hp = self.hyper_parameter
        self.model.to(self.device)
        self.model.train()
        try:
            self.exec_hooks(ExecutorHookPoint.BEFORE_EXECUTE, executor=self)
            optimizer = self.get_optimizer()
            steps = hp.epoch * max(math.ceil(self.dataset_size / hp.batch_size), 1)
            scheduler = torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=steps)
            for epoch in range(1, hp.epoch + 1):
                start = time.monotonic()
                acc = torch.zeros(3, dtype=torch.float64, device=self.device)
                for batch_index, (x, y) in enumerate(self._batches(self.phase, hp.batch_size, shuffle=True)):
                    optimizer.zero_grad(set_to_none=True)
                    with self._autocast():
                        out = self.model(x)
                        loss = self.loss_fun(out.float(), y)
                    loss.backward()
                    if self.has_hook(ExecutorHookPoint.OPTIMIZER_STEP):
                        self.exec_hooks(ExecutorHookPoint.OPTIMIZER_STEP, executor=self)
                    else:
                        optimizer.step()
                    scheduler.step()
                    acc[0] += loss.detach().double() * y.shape[0]
                    acc[1] += (out.argmax(-1) == y).sum()
                    acc[2] += y.shape[0]
                    self.exec_hooks(ExecutorHookPoint.AFTER_BATCH, executor=self, epoch=epoch, batch_index=batch_index)
                acc = acc.cpu()
                n = max(float(acc[2]), 1.0)
                self.performance_metric.set(epoch, float(acc[0]) / n, float(acc[1]) / n)
                get_logger().info("%sepoch: %d, train in %.3f s, loss: %e, accuracy: %.2f%%", self._prefix, epoch,
                                  time.monotonic() - start, float(acc[0]) / n, 100 * float(acc[1]) / n)
                if self.dataset_collection.has_dataset(MachineLearningPhase.Validation) and \
                        self.has_hook(ExecutorHookPoint.AFTER_VALIDATION):
                    self.model.eval()
                    vm = self._evaluate(MachineLearningPhase.Validation)
                    self.model.train()
                    self.exec_hooks(ExecutorHookPoint.AFTER_VALIDATION, executor=self, epoch=epoch,
                                    validation_metric=vm)
                self.exec_hooks(ExecutorHookPoint.AFTER_EPOCH, executor=self, epoch=epoch)
            self.exec_hooks(ExecutorHookPoint.AFTER_EXECUTE, executor=self, epoch=hp.epoch)
        except StopExecutingException:
            pass


class Inferencer(Executor):
    def inference(self, **kwargs) -> None:
        start = time.monotonic()
        self.model.eval()
        metric = self._evaluate(self.phase)
        self.performance_metric.set(1, metric["loss"], metric["accuracy"])
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        timing = os.environ.get("REF_TIMING_FILE")
        if timing and self.hook_config.log_performance_metric:  # round-level evaluations only (not Shapley subsets)
            # the benchmark harness brackets rounds with the completion time of each evaluation
            with open(timing, "at", encoding="utf8") as f:
                f.write(f"{time.time()}\n")
        if self.hook_config.log_performance_metric:
            get_logger().info("%s%s in %.3f s, loss: %e, accuracy: %.2f%%", self._prefix, self.phase.value,
                              time.monotonic() - start, metric["loss"], 100 * metric["accuracy"])
