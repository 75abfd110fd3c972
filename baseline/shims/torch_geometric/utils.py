import torch


def mask_to_index(mask):
    return mask.nonzero(as_tuple=False).view(-1)


def index_to_mask(index, size=None):
    size = int(index.max()) + 1 if size is None else size
    mask = torch.zeros(size, dtype=torch.bool)
    mask[index] = True
    return mask


def coalesce(edge_index, num_nodes=None):
    n = int(edge_index.max()) + 1 if num_nodes is None else num_nodes
    key = torch.unique(edge_index[0] * n + edge_index[1])
    return torch.stack([key // n, key % n])
