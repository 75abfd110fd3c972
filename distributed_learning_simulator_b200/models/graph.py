"""Own GCN layers and the three graph models the reference configs name
(``TwoGCN``, ``OneGCN``, ``SimpleGCN``; /root/reference/conf/fed_gnn/*.yaml,
conf/fed_aas/*.yaml).  ``torch_geometric`` is not installable here, so
``MessagePassing`` is a minimal base class that exists for the one purpose the
reference needs it for: locating the layers whose inputs are hooked for
boundary-embedding exchange (worker/graph_worker.py:375-380)."""
from typing import Any

import torch
from torch import nn

from ..ops import graph as graph_ops
from .registry import register_model


class MessagePassing(nn.Module):
    """Base class of layers called as ``layer(x, edge_index, ...)``."""


def gcn_norm(edge_index: torch.Tensor, num_nodes: int, dtype: torch.dtype) -> tuple[torch.Tensor, torch.Tensor]:
    """Add self loops, return (edge_index, D^-1/2 A D^-1/2 edge weights)."""
    loop = torch.arange(num_nodes, device=edge_index.device)
    ei = torch.cat([edge_index, torch.stack([loop, loop])], dim=1)
    deg = torch.zeros(num_nodes, dtype=dtype, device=ei.device).index_add_(
        0, ei[1], torch.ones(ei.shape[1], dtype=dtype, device=ei.device))
    dinv = deg.clamp(min=1.0).rsqrt()
    return ei, dinv[ei[0]] * dinv[ei[1]]


class GCNConv(MessagePassing):
    def __init__(self, in_channels: int, out_channels: int, bias: bool = True) -> None:
        super().__init__()
        self.lin = nn.Linear(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor, *args: Any, **kwargs: Any) -> torch.Tensor:
        h = self.lin(x)
        # normalisation + neighbour sum + bias: csrc/graph.cu on a GPU, the index_add_ oracle on CPU
        return graph_ops.gcn_aggregate(h, edge_index, self.bias)


class _GCN(nn.Module):
    def __init__(self, in_features: int, num_classes: int, hidden: list[int], dropout: float = 0.5) -> None:
        super().__init__()
        dims = [in_features, *hidden, num_classes]
        self.convs = nn.ModuleList(GCNConv(dims[i], dims[i + 1]) for i in range(len(dims) - 1))
        self.dropout = dropout

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor, n_id: torch.Tensor | None = None,
                **kwargs: Any) -> torch.Tensor:
        for i, conv in enumerate(self.convs):
            x = conv(x, edge_index)
            if i != len(self.convs) - 1:
                x = torch.relu(x)
                x = torch.nn.functional.dropout(x, p=self.dropout, training=self.training)
        return x


def _features(dc: Any) -> int:
    return int(dc.spec["features"])


def _classes(dc: Any) -> int:
    return int(dc.spec["classes"])


register_model("SimpleGCN", lambda dataset_collection, **kw: _GCN(_features(dataset_collection), _classes(dataset_collection), [], 0.0))
register_model("OneGCN", lambda dataset_collection, **kw: _GCN(_features(dataset_collection), _classes(dataset_collection), [kw.get("hidden", 64)]))
register_model("TwoGCN", lambda dataset_collection, **kw: _GCN(_features(dataset_collection), _classes(dataset_collection), [kw.get("hidden", 64)] * 2))
