import torch
from torch import nn


class ModelUtil:
    def __init__(self, model: nn.Module) -> None:
        self.model = model

    def get_parameter_dict(self, detach: bool = True) -> dict:
        return {k: (v.detach() if detach else v) for k, v in self.model.named_parameters()}

    def get_parameter_list(self, detach: bool = True) -> torch.Tensor:
        return torch.cat([p.detach().reshape(-1) for p in self.model.parameters()])

    def load_parameter_dict(self, parameter_dict: dict, **kwargs) -> None:
        own = dict(self.model.named_parameters())
        with torch.no_grad():
            for k, v in parameter_dict.items():
                own[k].copy_(v.to(own[k].device))

    def get_modules(self) -> list:
        return list(self.model.named_modules())

    def get_sub_modules(self) -> list:
        return [(n, m) for n, m in self.model.named_modules() if not list(m.children())]

    def get_module_blocks(self, block_types) -> list:
        def matches(m, t):
            return type(m).__name__ == t if isinstance(t, str) else isinstance(m, t)

        leaves = self.get_sub_modules()
        types = sorted(block_types, key=len, reverse=True)
        blocks, i = [], 0
        while i < len(leaves):
            for t in types:
                if i + len(t) <= len(leaves) and all(matches(leaves[i + j][1], t[j]) for j in range(len(t))):
                    blocks.append(leaves[i: i + len(t)])
                    i += len(t)
                    break
            else:
                i += 1
        return blocks

    def disable_running_stats(self) -> None:
        for m in self.model.modules():
            if isinstance(m, nn.modules.batchnorm._BatchNorm):
                m.track_running_stats = False
                m.register_buffer("running_mean", None)
                m.register_buffer("running_var", None)
                m.register_buffer("num_batches_tracked", None)
