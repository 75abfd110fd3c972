"""Front-end of the hand-written tcgen05 GEMM (csrc/gemm_tcgen05.cu).

``linear_bf16(x, weight, bias)`` computes ``act(x @ weight.T + bias)`` with bf16 operands and
fp32 accumulation in tensor memory; a leading batch dimension on ``x`` (and optionally on
``weight``/``bias``) runs one independent GEMM per entry in the same launch -- the Shapley
coalition evaluator uses that with batch = coalition."""
import torch

from . import runtime


def available() -> bool:
    return runtime.available()


def linear_bf16(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False,
                out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """x: [M, K] or [B, M, K]; weight: [N, K] or [B, N, K]; bias: [N] or [B, N]."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    squeeze = x.dim() == 2
    if squeeze:
        x = x.unsqueeze(0)
    batch, m, k = x.shape
    shared = weight.dim() == 2
    n = weight.shape[-2]
    assert weight.shape[-1] == k and (shared or weight.shape[0] == batch)
    if k % 8:  # TMA rows must be 16-byte multiples: pad K with zeros (contributes nothing)
        pad = 8 - k % 8
        x = torch.nn.functional.pad(x, (0, pad))
        weight = torch.nn.functional.pad(weight, (0, pad))
        k += pad
    x = x.contiguous()
    weight = weight.contiguous()
    if bias is not None:
        bias = bias.float().contiguous()
        assert bias.shape[-1] == n and (bias.dim() == 1) == shared
    out = torch.empty((batch, m, n), dtype=out_dtype, device=x.device)
    runtime.call("dls_gemm_bf16_tn", x.data_ptr(), weight.data_ptr(), runtime.ptr(bias), out.data_ptr(),
                 0 if out_dtype == torch.float32 else 1, m, n, k, batch, int(shared), int(relu),
                 runtime.stream_ptr(x.device))
    return out[0] if squeeze else out


def oracle_linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
                  relu: bool = False) -> torch.Tensor:
    """Plain PyTorch fp32 statement of the same op on the same (bf16-rounded) operands."""
    out = torch.matmul(x.float(), weight.float().transpose(-1, -2))
    if bias is not None:
        out = out + (bias.float() if bias.dim() == 1 else bias.float().unsqueeze(-2))
    return torch.relu(out) if relu else out


def subset_linear(x: torch.Tensor, worker_flats: list[torch.Tensor], pointer_table: torch.Tensor,
                  coefs: torch.Tensor, w_offset: int, bias_offset: int | None, n_layer: int, k_w: int,
                  stacked: bool = False, relu: bool = False, out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Coalition GEMM (csrc/subset_gemm.cu): ``act(x[s] @ W_s.T + b_s)`` with
    ``W_s = sum_i coefs[s, i] * worker_flats[i][w_offset:...].view(n_layer, k_w)`` formed inside the
    kernel.  x: bf16 [S, M, K] (or [M, K], shared by all coalitions); K may exceed k_w (zero padding
    up to a multiple of 8).  Returns [S, M, n_layer], or [M, S * n_layer] when ``stacked``."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and coefs.dtype == torch.float32 and coefs.is_contiguous()
    n_sets, n_workers = coefs.shape
    assert n_workers == len(worker_flats) == pointer_table.numel()
    shared = x.dim() == 2
    if shared:
        x = x.unsqueeze(0)
    assert stacked <= shared and (shared or x.shape[0] == n_sets)
    m, k_a = x.shape[1], x.shape[2]
    assert k_a >= k_w
    if k_a % 8:
        x = torch.nn.functional.pad(x, (0, 8 - k_a % 8))
        k_a = x.shape[2]
    x = x.contiguous()
    aligned = all(f.data_ptr() % 16 == 0 for f in worker_flats)
    shape = (m, n_sets * n_layer) if stacked else (n_sets, m, n_layer)
    out = torch.empty(shape, dtype=out_dtype, device=x.device)
    runtime.call("dls_subset_gemm_bf16", x.data_ptr(), int(shared), pointer_table.data_ptr(), n_workers,
                 int(w_offset), -1 if bias_offset is None else int(bias_offset), coefs.data_ptr(), n_sets,
                 n_layer, int(stacked), out.data_ptr(), 0 if out_dtype == torch.float32 else 1, m, k_a, k_w,
                 int(relu), int(aligned), runtime.stream_ptr(x.device))
    return out


def oracle_subset_linear(x: torch.Tensor, worker_flats: list[torch.Tensor], coefs: torch.Tensor, w_offset: int,
                         bias_offset: int | None, n_layer: int, k_w: int, stacked: bool = False,
                         relu: bool = False) -> torch.Tensor:
    """fp32 statement of ``subset_linear`` on the same bf16-rounded operands."""
    stack = torch.stack([f[w_offset: w_offset + n_layer * k_w].view(n_layer, k_w) for f in worker_flats])
    w = torch.einsum("si,ink->snk", coefs, stack).to(torch.bfloat16).float()       # [S, n, k]
    xs = x.float()[..., :k_w]
    if xs.dim() == 2:
        xs = xs.unsqueeze(0).expand(coefs.shape[0], -1, -1)
    out = torch.matmul(xs, w.transpose(1, 2))                                       # [S, M, n]
    if bias_offset is not None:
        b = torch.stack([f[bias_offset: bias_offset + n_layer] for f in worker_flats])
        out = out + (coefs @ b).unsqueeze(1)
    if relu:
        out = torch.relu(out)
    if stacked:
        out = out.permute(1, 0, 2).reshape(out.shape[1], -1)
    return out


STATS = {"linear_calls": 0}


def install_inference_linear(model: torch.nn.Module) -> int:
    """Route every ``nn.Linear`` of an evaluation-only model through the tcgen05 GEMM.

    Applies when the call runs without autograd on CUDA (the server's test pass); training calls
    and CPU calls keep the stock path.  bf16 operands, fp32 accumulation and fp32 logits -- the
    same precision class as the bf16 autocast the evaluation otherwise runs under.
    Returns the number of layers patched."""
    import types

    def forward(self: torch.nn.Linear, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and not torch.is_grad_enabled() and available() and x.dim() >= 2 \
                and x.dtype in (torch.bfloat16, torch.float32) and torch.is_autocast_enabled():
            lead = x.shape[:-1]
            out = linear_bf16(x.reshape(-1, x.shape[-1]).to(torch.bfloat16), self.weight.to(torch.bfloat16), self.bias)
            STATS["linear_calls"] += 1
            return out.view(*lead, out.shape[-1])
        return torch.nn.functional.linear(x, self.weight, self.bias)

    patched = 0
    for module in model.modules():
        if type(module) is torch.nn.Linear and not getattr(module, "_dls_tcgen05", False):
            module.forward = types.MethodType(forward, module)
            module._dls_tcgen05 = True
            patched += 1
    return patched
