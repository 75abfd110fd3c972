"""signSGD worker: every optimizer step uploads one bit per parameter (32 signs per word,
``ops.misc.sign_pack``) and applies the server's majority vote with the local learning
rate and momentum.  At the end of each epoch worker 0 ships its parameters so the server
can log ``test loss`` / accuracy (the offline scraper keys on "test loss":
reference analysis/analyze_log.py:23-29)."""
from typing import Any

import torch

from ...engine.ml_type import ExecutorHookPoint
from ...message import Message, ParameterMessage
from ...ops import misc as misc_ops
from ...worker.gradient_worker import GradientWorker


class SignSGDWorker(GradientWorker):
    def _before_training(self) -> None:
        first = not self.trainer.has_named_hook("step")
        super()._before_training()
        if first:
            self.trainer.dataset_collection.remove_dataset(self._test_phase())
            self.trainer.append_named_hook(ExecutorHookPoint.AFTER_EPOCH, "report_epoch", self._report_epoch)

    @staticmethod
    def _test_phase() -> Any:
        from ...engine.ml_type import MachineLearningPhase

        return MachineLearningPhase.Test

    def _process_gradient(self, gradient: torch.Tensor) -> Any:
        return misc_ops.sign_pack(gradient)

    def _decode_gradient(self, payload: Any, numel: int) -> torch.Tensor:
        if isinstance(payload, dict) and "sign_rows" in payload:  # no fused path here (CPU / non-flat model)
            return misc_ops.sign_majority([misc_ops.as_local_words(r) for r in payload["sign_rows"]], numel)
        return payload

    def _apply_fused(self, result: Message, flat: torch.Tensor, momentum_buf: torch.Tensor, lr: float, hp: Any,
                     momentum_ready: bool) -> bool:
        """Vote + update in one launch over every worker's packed words (peer-mapped for other ranks)."""
        payload = result.other_data.get("gradient")
        if not (isinstance(payload, dict) and "sign_rows" in payload):
            return False
        util = self.trainer.model_util
        misc_ops.sign_vote_sgd(payload["sign_rows"], flat.numel(), flat, momentum_buf if hp.momentum else None, lr,
                               hp.momentum, 0.0, False, momentum_ready, shadow=util.shadow)
        return True

    def _report_epoch(self, **kwargs: Any) -> None:
        epoch = kwargs["epoch"]
        if self.worker_id == 0:
            util = self.trainer.model_util
            flat = util.flat_parameter
            from ...engine.model_util import FlatParameter

            parameter = FlatParameter(flat.detach().clone(), util.layout) if flat is not None else {
                k: v.clone() for k, v in util.get_parameter_dict().items()}
            self.send_data_to_server(ParameterMessage(in_round=True, parameter=parameter,
                                                      dataset_size=self.trainer.dataset_size,
                                                      other_data={"epoch": epoch}))
        else:
            self.send_data_to_server(Message(in_round=True, other_data={"epoch": epoch}))
        self._get_data_from_server()
