"""TransformerClassificationModel: embedding + sinusoidal positions + N encoder layers
+ masked mean pool + linear head.  Hyper-parameters come from ``model_kwargs`` in the
reference's IMDB configs (/root/reference/conf/fed_avg/imdb.yaml:16-21: max_len 300,
word_vector_name glove.6B.100d, num_encoder_layer 2, d_model 100, nhead 5).  GloVe
vectors cannot be downloaded here, so the embedding is randomly initialised with the
same width."""
import math
from typing import Any

import torch
from torch import nn

from .registry import register_model


class TransformerClassificationModel(nn.Module):
    def __init__(self, vocab_size: int, num_classes: int, max_len: int = 300, d_model: int = 100, nhead: int = 5,
                 num_encoder_layer: int = 2, dim_feedforward: int = 2048, dropout: float = 0.1,
                 word_vector_name: str | None = None, **kwargs: Any) -> None:
        super().__init__()
        self.d_model = d_model
        self.embedding = nn.Embedding(vocab_size, d_model, padding_idx=0)
        pe = torch.zeros(max_len, d_model)
        position = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
        pe[:, 0::2] = torch.sin(position * div_term)
        pe[:, 1::2] = torch.cos(position * div_term)[:, : d_model // 2]
        self.register_buffer("positional_encoding", pe.unsqueeze(0), persistent=False)
        self.encoder_layers = nn.ModuleList(
            nn.TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, batch_first=True)
            for _ in range(num_encoder_layer)
        )
        self.head = nn.Linear(d_model, num_classes)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        pad = tokens == 0
        x = self.embedding(tokens) * math.sqrt(self.d_model)
        x = x + self.positional_encoding[:, : x.shape[1]].to(x.dtype)
        for layer in self.encoder_layers:
            x = layer(x, src_key_padding_mask=pad)
        keep = (~pad).unsqueeze(-1).to(x.dtype)
        pooled = (x * keep).sum(dim=1) / keep.sum(dim=1).clamp(min=1.0)
        return self.head(pooled)


register_model(
    "TransformerClassificationModel",
    lambda dataset_collection, **kw: TransformerClassificationModel(
        vocab_size=int(dataset_collection.spec["vocab"]), num_classes=int(dataset_collection.spec["classes"]), **kw
    ),
)
