"""Per-epoch loss/accuracy bookkeeping with device-side accumulation.

Batch statistics are summed into device tensors; the host reads them once per epoch, so
a training epoch enqueues no synchronising call.  Mirrors what the reference reads from
``performance_metric`` (``get_epoch_metrics(1)``, ``get_loss``, ``get_accuracy``:
/root/reference/simulation_lib/server/server.py:52, worker/gradient_worker.py:122-127).
"""
import torch


class PerformanceMetric:
    d2h_bytes = 0  # process-wide count of metric bytes read back from the device

    def __init__(self) -> None:
        self._epoch_metrics: dict[int, dict[str, torch.Tensor]] = {}
        self._acc: torch.Tensor | None = None  # [loss_sum, correct, count]

    def start_epoch(self, device: torch.device) -> None:
        self._acc = torch.zeros(3, dtype=torch.float64, device=device)

    @property
    def accumulator(self) -> torch.Tensor:
        assert self._acc is not None
        return self._acc

    def add_batch(self, loss: torch.Tensor, output: torch.Tensor, targets: torch.Tensor) -> None:
        """``loss`` is the mean loss of the batch."""
        assert self._acc is not None
        n = targets.shape[0]
        correct = (output.argmax(dim=-1) == targets).sum()
        self._acc += torch.stack([loss.detach().double() * n, correct.double(),
                                  torch.tensor(float(n), dtype=torch.float64, device=loss.device)])

    def add_counts(self, loss_sum: torch.Tensor, correct: torch.Tensor, count: torch.Tensor | float) -> None:
        assert self._acc is not None
        self._acc[0] += loss_sum.double()
        self._acc[1] += correct.double()
        self._acc[2] += count if not isinstance(count, torch.Tensor) else count.double()

    def end_epoch(self, epoch: int) -> dict[str, torch.Tensor]:
        assert self._acc is not None
        acc = self._acc.cpu()
        if self._acc.is_cuda:
            PerformanceMetric.d2h_bytes += acc.numel() * acc.element_size()
        cnt = acc[2].clamp(min=1.0)
        m = {"loss": (acc[0] / cnt).float(), "accuracy": (acc[1] / cnt).float(), "sample_count": acc[2].float()}
        self._epoch_metrics[epoch] = m
        return m

    def set_epoch_metrics(self, epoch: int, loss: float, accuracy: float, count: float) -> None:
        self._epoch_metrics[epoch] = {"loss": torch.tensor(loss), "accuracy": torch.tensor(accuracy),
                                      "sample_count": torch.tensor(count)}

    def get_epoch_metrics(self, epoch: int) -> dict[str, float]:
        m = self._epoch_metrics[epoch]
        return {"loss": float(m["loss"]), "accuracy": float(m["accuracy"])}

    def get_loss(self, epoch: int) -> torch.Tensor:
        return self._epoch_metrics[epoch]["loss"]

    def get_accuracy(self, epoch: int) -> torch.Tensor:
        return self._epoch_metrics[epoch]["accuracy"]

    def clear(self) -> None:
        self._epoch_metrics.clear()
