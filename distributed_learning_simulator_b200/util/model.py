"""Loading a received global model into a trainer.

Same contract as /root/reference/simulation_lib/util/model.py:6-23: with
``reuse_learning_rate`` the optimizer's hyper-parameters (current LR, schedule position)
survive and only its state is cleared; otherwise optimizer and scheduler start fresh.
Running BatchNorm statistics are always disabled (model.py:23).
"""
from ..engine.trainer import Trainer
from ..engine.tensor import TensorDict
from ..message import as_flat


def load_parameters(trainer: Trainer, parameter_dict: TensorDict, reuse_learning_rate: bool) -> None:
    flat = as_flat(parameter_dict)
    if flat is not None and trainer.model_util.flat_parameter is not None \
            and flat.layout == trainer.model_util.layout:
        trainer.load_parameter_list(flat.flat, reuse_learning_rate=reuse_learning_rate)
    elif reuse_learning_rate:
        trainer.model_util.load_parameter_dict(parameter_dict)
        if trainer._optimizer is not None:  # noqa: SLF001 -- same package
            trainer._optimizer.state.clear()
        trainer.load_parameter_list(trainer.model_util.get_parameter_list(), reuse_learning_rate=True)
    else:
        trainer.load_parameter_dict(parameter_dict)
    trainer.model_util.disable_running_stats()
