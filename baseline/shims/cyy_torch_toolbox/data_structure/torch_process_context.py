import torch.multiprocessing


class TorchProcessContext:
    def get_ctx(self):
        return torch.multiprocessing.get_context("spawn")
