#!/usr/bin/env python3
"""One launch (after one warm-up) of the kernels added since the first ncu capture: CTA-pair GEMM, coalition
GEMM (weight-stationary cluster variant), grouped im2col, GCN aggregation, edge compaction, bf16-gradient SGD.
Workload for `ncu --set full`; prints nothing a benchmark would quote."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200.engine.model_util import ParameterLayout  # noqa: E402
from distributed_learning_simulator_b200.models.vision import ResNet18  # noqa: E402
from distributed_learning_simulator_b200.ops import gemm as G  # noqa: E402
from distributed_learning_simulator_b200.ops import graph as GR  # noqa: E402
from distributed_learning_simulator_b200.ops import misc as M  # noqa: E402
from distributed_learning_simulator_b200.ops import runtime as RT  # noqa: E402
from distributed_learning_simulator_b200.ops import sgd as S  # noqa: E402

dev = torch.device("cuda", 0)
RT.require()
a = torch.randn(8192, 4096, device=dev).bfloat16()
w = torch.randn(4096, 4096, device=dev).bfloat16()
NW, S_, Pl = 10, 32, 61706
arenas = [torch.randn(Pl + 6, device=dev)[:Pl] for _ in range(NW)]
table = RT.pointer_table(arenas, dev)
member = torch.rand(S_, NW) < 0.5
member[:, 0] |= ~member.any(dim=1)
coefs = (member.float() / member.float().sum(1, keepdim=True)).to(dev).contiguous()
xa = torch.randn(S_, 5000, 400, device=dev).bfloat16()
h = torch.randn(1024, 14, 14, 32 * 6, device=dev).bfloat16()
n_nodes, n_edges = 200000, 2000000
ei = torch.randint(0, n_nodes, (2, n_edges), device=dev)
feat = torch.randn(n_nodes, 64, device=dev)
mask = torch.rand(n_nodes, device=dev) < 0.6
model = ResNet18(3, 10)
layout = ParameterLayout.from_dict(dict(model.named_parameters()))
P = layout.total
flat, mom, shadow = torch.randn(P, device=dev), torch.zeros(P, device=dev), torch.zeros(P, device=dev).bfloat16()
flags_host = [int(len(s) >= 2 and o % 8 == 0) for s, o in zip(layout.shapes, layout.offsets)]
grads = [torch.randn(n, device=dev).to(torch.bfloat16 if f else torch.float32) for n, f in zip(layout.numels, flags_host)]
gtable = torch.tensor([g.data_ptr() for g in grads], dtype=torch.int64, device=dev)
flags = torch.tensor(flags_host, dtype=torch.uint8, device=dev)
offsets = layout.segment_offsets(dev)
lr = torch.full((8,), 0.01, device=dev)
state = torch.tensor([0, 1, 0, 0], dtype=torch.int32, device=dev)
for _ in range(2):
    G.linear_bf16(a, w, None, out_dtype=torch.bfloat16)
    G.subset_linear(xa, arenas, table, coefs, 2572, 2572 + 120 * 400, 120, 400, relu=True, out_dtype=torch.bfloat16)
    M.im2col_grouped(h, 32, 5, True)
    GR.gcn_aggregate(feat, ei, None)
    GR.filter_edges(ei, mask)
    S.sgd_step(flat, mom, gtable, offsets, len(layout.names), lr, state, 1e-4, 0.9, grad_is_bf16=flags, shadow=shadow)
    torch.cuda.synchronize()
print("done")
