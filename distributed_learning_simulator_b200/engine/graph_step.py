"""CUDA-graph training step over the flat parameter arena (the B200 training path).

One optimizer step = [gather batch] -> forward -> loss -> backward -> [fused SGD] ->
[tick] -> [metric accumulate]; the bracketed stages are this framework's kernels
(csrc/sgd.cu), the rest is PyTorch autograd under bf16 autocast when ``use_amp`` is on.
The whole sequence is captured ONCE per batch size and replayed for every step of every
round: the step's batch index, learning rate and momentum-initialised flag all live in
device memory (``state``/``lr_table``), so nothing captured depends on a host value and
the host only enqueues ``replay()`` -- no per-step Python dispatch of ~200 kernels, no
synchronisation until the epoch's metrics are read.

The reference's counterpart is the external ``Trainer.train()`` eager loop with
``torch.optim.SGD`` (one ATen launch per tensor per step) invoked at
/root/reference/simulation_lib/worker/worker.py:86-88.
"""
import os
import time
from typing import Any

import torch

from ..ops import conv as conv_ops
from ..ops import runtime
from ..ops import sgd as sgd_ops
from .dataloader import SampleLoader
from .device import capture_guard, capture_stream
from .ml_type import ExecutorHookPoint, MachineLearningPhase


# captures are serialised and exclude other threads' cuDNN/cuBLAS eager sections (device.CaptureGuard);
# graph replays and this framework's own kernels keep running (thread-local capture mode)
_capture_lock = capture_guard.exclusive

# process-wide counters read by bench.py
REPLAY_COUNTER = [0]
H2D_BYTES = [0]
OWN_REPLAYED_KERNELS = [0]  # hand-written kernels executed through graph replays
CAPTURE_LAUNCHES = [0]      # launch-counter increments that happened while warming up / capturing


def _same_physical_layout(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Equal strides on every dimension that has more than one element."""
    return a.shape == b.shape and all(sa == sb for n, sa, sb in zip(a.shape, a.stride(), b.stride()) if n > 1)


class _CapturedStep:
    def __init__(self) -> None:
        self.graph: torch.cuda.CUDAGraph | None = None
        self.x: torch.Tensor | None = None
        self.y: torch.Tensor | None = None
        self.grad_table: torch.Tensor | None = None
        self.grads: list = []
        self.own_kernels = 0


class GraphTrainStep:
    def __init__(self, trainer: Any) -> None:
        runtime.require()
        self.trainer = trainer
        self.device = trainer.device
        self.util = trainer.model_util
        self.flat = self.util.flatten_parameters(with_grad=False)
        self.layout = self.util.layout
        self.n_tensors = len(self.layout.names)
        self.offsets = torch.tensor(self.layout.offsets, dtype=torch.int64, device=self.device)
        self.momentum_buf = torch.zeros_like(self.flat)
        # mixed precision: matrices/filters of the model are bf16 views of a shadow arena that the
        # fused SGD kernel rewrites, so no weight or gradient cast kernel runs inside the step
        self.use_shadow = bool(trainer.hook_config.use_amp) and os.environ.get("DLS_BF16_SHADOW", "1") != "0"
        self.shadow: torch.Tensor | None = None
        self.shadow_flags: torch.Tensor | None = None
        if self.use_shadow:
            self.shadow_flags = torch.tensor(self.util.shadow_flags(), dtype=torch.uint8, device=self.device)
        # convolutions run on this framework's kernels (ops/conv.py); their weight-gradient kernels go to a
        # side stream -- a parallel branch of the captured graph -- and write fp32 sinks that the fused SGD
        # kernel reads in place
        self.side_stream: torch.cuda.Stream | None = None
        self.wgrad_sinks: dict[int, list] = {}
        self._sink_of_param: dict[int, list] = {}
        n_convs = conv_ops.install(trainer.model)
        if n_convs > 0 and os.environ.get("DLS_SPLIT_BACKWARD", "1") != "0":
            self.side_stream = torch.cuda.Stream(device=self.device)
        self.grad_flags: torch.Tensor | None = None
        # per step: [2][Cout] for every convolution's statistics epilogue + [2][C] for every BatchNorm's backward sums
        stats_floats = sum(2 * m.out_channels for m in trainer.model.modules() if type(m) is torch.nn.Conv2d) + \
            sum(2 * m.num_features for m in trainer.model.modules() if isinstance(m, torch.nn.BatchNorm2d))
        self.stats_buf = torch.zeros(stats_floats, dtype=torch.float32, device=self.device) if stats_floats else None
        # state: [0] lr/optimizer step, [1] momentum initialised, [2] batch index in epoch
        self.state = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.acc = torch.zeros(3, dtype=torch.float64, device=self.device)
        self.lr_table = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.order: torch.Tensor | None = None
        self.steps: dict[int, _CapturedStep] = {}
        self.fresh = True
        self.active = False
        self._lr_host: torch.Tensor | None = None
        self._host_step = 0
        self.replays = 0

    # ---- optimizer state -------------------------------------------------------------
    def reset_optimizer_state(self, keep_schedule: bool = False) -> None:
        if keep_schedule:
            with torch.cuda.stream(self.trainer.stream):
                self.state[1:2].zero_()
        else:
            self.fresh = True

    def current_lr(self) -> float:
        if self._lr_host is None:
            return self.trainer.hyper_parameter.learning_rate
        return float(self._lr_host[min(self._host_step, self._lr_host.numel() - 1)])

    # ---- capture -----------------------------------------------------------------------
    def _one_step(self, step: _CapturedStep, x_all: torch.Tensor, y_all: torch.Tensor, n_order: int) -> None:
        t = self.trainer
        hp = t.hyper_parameter
        sgd_ops.gather_batch(x_all, y_all, self.order, n_order, self.state, 2, step.x, step.y,
                             batch_stride=hp.batch_size)
        if self.stats_buf is not None:
            self.stats_buf.zero_()  # one fill for the BatchNorm statistics of every convolution epilogue
        with t._autocast(), conv_ops.split_backward(self.side_stream, self.wgrad_sinks), \
                conv_ops.stats_arena(self.stats_buf):
            output = t.model(step.x)
            loss = t.loss_fun(output.float(), step.y)
        loss.backward()
        if self.side_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.side_stream)  # join before the optimizer
        if step.grad_table is None:
            step.grad_table = torch.zeros(self.n_tensors, dtype=torch.int64, device=self.device)
        sgd_ops.sgd_step(self.flat, self.momentum_buf if hp.momentum else None, step.grad_table, self.offsets,
                         self.n_tensors, self.lr_table, self.state, hp.weight_decay, hp.momentum,
                         grad_is_bf16=self.grad_flags, shadow=self.shadow)
        sgd_ops.sgd_tick(self.state)
        sgd_ops.metric_accumulate(output.float().contiguous(), step.y, loss.detach().reshape(1), self.acc)

    def _make_sinks(self) -> None:
        """One persistent fp32 buffer per convolution filter, in the filter's physical layout."""
        if self.wgrad_sinks or os.environ.get("DLS_WGRAD_SINK", "1") == "0":
            return
        for module in self.trainer.model.modules():
            if type(module) is torch.nn.Conv2d and getattr(module, "_dls_own_conv", False):
                w = module.weight
                if not w.requires_grad:
                    continue
                # [buffer, accumulate?]: both are decided by the layer's first forward (ops/conv.py)
                entry = [None, None]
                self.wgrad_sinks[id(module)] = entry
                self._sink_of_param[id(w)] = entry

    def _fill_grad_table(self, step: _CapturedStep) -> None:
        params = list(self.trainer.model.parameters())
        step.grads = [p.grad for p in params]
        ptrs, flags = [], []
        base_flags = self.util.shadow_flags() if self.use_shadow else [0] * len(params)
        for p, g, f in zip(params, step.grads, base_flags):
            if g is None and id(p) in self._sink_of_param and self._sink_of_param[id(p)][0] is not None:
                # the convolution wrote its weight gradient into the sink (fp32, accumulating): bit 2 lets
                # the update clear it for the next step; bit 1 keeps the bf16 shadow rewrite
                sink, accumulate = self._sink_of_param[id(p)]
                ptrs.append(sink.data_ptr())
                if sink.dtype == torch.bfloat16:
                    flags.append(1)  # stored once per step in the shadow's dtype: an ordinary bf16 gradient
                else:
                    flags.append((2 if f else 0) | (4 if accumulate else 0))
                continue
            # the fused update walks physical memory: a gradient must be laid out like its parameter
            # (autograd's layout contract gives channels-last grads for channels-last weights)
            assert g is None or (g.dtype == p.dtype and _same_physical_layout(g, p)), (g.stride(), p.stride())
            ptrs.append(0 if g is None else g.data_ptr())
            flags.append(f)
        step.grad_table.copy_(torch.tensor(ptrs, dtype=torch.int64))
        new_flags = torch.tensor(flags, dtype=torch.uint8)
        if self.grad_flags is None:
            self.grad_flags = new_flags.to(self.device)
        else:
            self.grad_flags.copy_(new_flags)

    def _capture(self, batch: int, x_all: torch.Tensor, y_all: torch.Tensor, n_order: int) -> _CapturedStep:
        with _capture_lock():
            before = runtime.launch_count()
            step = self._capture_locked(batch, x_all, y_all, n_order)
            CAPTURE_LAUNCHES[0] += runtime.launch_count() - before
            return step

    def _capture_locked(self, batch: int, x_all: torch.Tensor, y_all: torch.Tensor, n_order: int) -> _CapturedStep:
        t = self.trainer
        assert self.trainer.hyper_parameter.optimizer_name.lower() == "sgd", "the fused step implements SGD"
        step = _CapturedStep()
        step.x = torch.empty((batch, *x_all.shape[1:]), dtype=x_all.dtype, device=self.device)
        if step.x.dim() == 4 and step.x.shape[1] > 1 and any(f for f in self.layout.formats):
            step.x = step.x.contiguous(memory_format=torch.channels_last)  # match the channels-last weights
        step.y = torch.empty(batch, dtype=torch.int64, device=self.device)
        params = list(t.model.parameters())
        self._make_sinks()
        if self.grad_flags is None:
            flags = self.util.shadow_flags() if self.use_shadow else [0] * len(params)
            self.grad_flags = torch.tensor(flags, dtype=torch.uint8, device=self.device)
        # warm-up (lazy init, library heuristics) must not change the training state
        saved = (self.flat.clone(), self.momentum_buf.clone(), self.state.clone(), self.acc.clone())
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                for p in params:
                    p.grad = None
                self._one_step(step, x_all, y_all, n_order)
                self._fill_grad_table(step)
        torch.cuda.current_stream(self.device).wait_stream(side)
        for p in params:
            p.grad = None
        torch.cuda.synchronize(self.device)
        # A capture can be invalidated by what another role thread does on the device at the same moment (seen
        # with dozens of workers per GPU); the step is idempotent up to the saved state, so it is simply retried.
        for attempt in range(4):
            graph = torch.cuda.CUDAGraph()
            own_before = runtime.launch_count()
            try:
                # the per-device capture stream (engine/device.py)
                with torch.cuda.graph(graph, stream=capture_stream(self.device), capture_error_mode="thread_local"):
                    self._one_step(step, x_all, y_all, n_order)
                break
            except RuntimeError as e:
                if attempt == 3:
                    raise
                from .log import get_logger

                get_logger().warning("training-step capture failed (%s), retrying", str(e).splitlines()[0])
                del graph
                for p in params:
                    p.grad = None
                torch.cuda.synchronize(self.device)
        step.own_kernels = runtime.launch_count() - own_before  # own kernel nodes in one replay
        self._fill_grad_table(step)
        self.flat.copy_(saved[0])
        self.util.refresh_shadow()
        self.momentum_buf.copy_(saved[1])
        self.state.copy_(saved[2])
        self.acc.copy_(saved[3])
        step.graph = graph
        torch.cuda.synchronize(self.device)
        return step

    # ---- run one train() -------------------------------------------------------------
    def run(self) -> None:
        t = self.trainer
        hp = t.hyper_parameter
        loader = t._loader(MachineLearningPhase.Training)
        assert isinstance(loader, SampleLoader)
        self.active = True
        batch = hp.batch_size
        with torch.cuda.stream(t.stream):
            if self.use_shadow:
                shadow = self.util.enable_shadow()
                if self.shadow is None or shadow.data_ptr() != self.shadow.data_ptr():
                    self.steps.clear()  # captured kernels hold the shadow's address
                self.shadow = shadow
                self.util.refresh_shadow()  # aggregation may have written the fp32 arena in place
            x_all, y_all, order0 = loader.epoch_tensors()  # shuffled order for epoch 1
            n = order0.numel()
            full, rem = divmod(n, batch)
            steps_per_epoch = full + (1 if rem else 0)
            total_steps = hp.epoch * steps_per_epoch
            if self.order is None or self.order.numel() < n:
                self.order = torch.zeros(max(n, 1), dtype=torch.int64, device=self.device)
                self.steps.clear()
            key_full, key_rem = (batch, x_all.data_ptr(), n), (rem, x_all.data_ptr(), n)
            if self.fresh:
                self._lr_host = hp.lr_table(total_steps)
                if self.lr_table.numel() < self._lr_host.numel():
                    # the captured kernels hold the table's address: size it before capturing
                    self.lr_table = torch.zeros(max(self._lr_host.numel(), 4096), dtype=torch.float32,
                                                device=self.device)
                    self.steps.clear()
            if full and key_full not in self.steps:
                self.steps[key_full] = self._capture(batch, x_all, y_all, n)
            if rem and key_rem not in self.steps:
                self.steps[key_rem] = self._capture(rem, x_all, y_all, n)
            if self.fresh:
                self.state.zero_()
                self._host_step = 0
                m = self._lr_host.numel()
                self.lr_table[:m].copy_(self._lr_host.to(self.device, non_blocking=True))
                self.lr_table[m:].fill_(float(self._lr_host[-1]))
                self.fresh = False
            for epoch in range(1, hp.epoch + 1):
                epoch_start = time.monotonic()
                t.exec_hooks(ExecutorHookPoint.BEFORE_EPOCH, executor=t, epoch=epoch)
                if epoch > 1:
                    # pinned-host storage re-copies the shard into the same staging buffers
                    order0 = loader.epoch_tensors()[2]
                self.order[:n].copy_(order0)
                self.state[2:3].zero_()
                self.acc.zero_()
                for _ in range(full):
                    self.steps[key_full].graph.replay()
                if rem:
                    self.steps[key_rem].graph.replay()
                self.replays += steps_per_epoch
                REPLAY_COUNTER[0] += steps_per_epoch
                OWN_REPLAYED_KERNELS[0] += full * self.steps[key_full].own_kernels if full else 0
                OWN_REPLAYED_KERNELS[0] += self.steps[key_rem].own_kernels if rem else 0
                self._host_step += steps_per_epoch
                t.performance_metric.start_epoch(self.device)
                t.performance_metric.accumulator.copy_(self.acc)
                t._epoch_end(epoch, epoch_start)
        H2D_BYTES[0] += loader.h2d_bytes
        self.active = False
