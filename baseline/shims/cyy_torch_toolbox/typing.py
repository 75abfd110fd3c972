import torch

TensorDict = dict[str, torch.Tensor]
