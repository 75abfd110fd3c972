"""Front-ends of the training-step kernels (csrc/sgd.cu): multi-tensor fused SGD with a
device-side LR schedule, batch gather driven by a device-side step counter, metric
accumulation without host synchronisation."""
import torch

from . import runtime


def sgd_step(params: torch.Tensor, momentum_buf: torch.Tensor | None, grad_table: torch.Tensor,
             tensor_offsets: torch.Tensor, n_tensors: int, lr_table: torch.Tensor, state: torch.Tensor,
             weight_decay: float, momentum: float, dampening: float = 0.0, nesterov: bool = False,
             grad_is_bf16: torch.Tensor | None = None, shadow: torch.Tensor | None = None) -> None:
    """``grad_is_bf16`` (uint8 [n_tensors]) + ``shadow`` (bf16 arena): mixed-precision mode, see
    ``ModelUtil.enable_shadow``."""
    runtime.call("dls_sgd_step", params.data_ptr(), runtime.ptr(momentum_buf), grad_table.data_ptr(),
                 tensor_offsets.data_ptr(), n_tensors, lr_table.data_ptr(), lr_table.numel(), state.data_ptr(),
                 float(weight_decay), float(momentum), float(dampening), int(nesterov),
                 runtime.ptr(grad_is_bf16), runtime.ptr(shadow), runtime.stream_ptr(params.device))


def sgd_tick(state: torch.Tensor) -> None:
    runtime.call("dls_sgd_tick", state.data_ptr(), runtime.stream_ptr(state.device))


def gather_batch(x_all: torch.Tensor, y_all: torch.Tensor, order: torch.Tensor, n_order: int, state: torch.Tensor,
                 state_index: int, x_out: torch.Tensor, y_out: torch.Tensor, batch_stride: int | None = None) -> None:
    """``x_out`` may be a channels-last 4-D tensor: rows are then written H,W,C."""
    batch = x_out.shape[0]
    assert x_out.dtype == x_all.dtype and x_out.element_size() % 4 == 0
    row = x_out[0].numel() * (x_out.element_size() // 4)  # rows are copied as 32-bit words (token ids: 2 per id)
    nhwc = 0
    if x_out.dim() == 4 and not x_out.is_contiguous():
        assert x_out.is_contiguous(memory_format=torch.channels_last) and x_out.element_size() == 4
        nhwc = int(x_out.shape[1])
    else:
        assert x_out.is_contiguous()
    runtime.call("dls_gather_batch", x_all.data_ptr(), y_all.data_ptr(), order.data_ptr(), int(n_order),
                 state.data_ptr(), int(state_index), int(batch), int(batch_stride or batch), int(row), nhwc,
                 x_out.data_ptr(), y_out.data_ptr(),
                 runtime.stream_ptr(x_out.device))


def metric_accumulate(logits: torch.Tensor, targets: torch.Tensor, loss: torch.Tensor, acc: torch.Tensor) -> None:
    assert logits.dtype == torch.float32 and logits.is_contiguous()
    runtime.call("dls_metric_accumulate", logits.data_ptr(), targets.data_ptr(), loss.data_ptr(),
                 int(logits.shape[0]), int(logits.shape[1]), acc.data_ptr(), runtime.stream_ptr(logits.device))


def oracle_sgd(params, mom, grads, lr, weight_decay, momentum, dampening, nesterov, mom_ready):
    """torch.optim.SGD semantics on flat vectors (test oracle)."""
    d = grads + weight_decay * params
    if momentum != 0:
        mom = momentum * mom + (1 - dampening) * d if mom_ready else d.clone()
        d = d + momentum * mom if nesterov else mom
    return params - lr * d, mom
