import torch


class MessagePassing(torch.nn.Module):
    pass
