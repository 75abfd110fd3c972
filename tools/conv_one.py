#!/usr/bin/env python3
"""Runs ONE convolution kernel a few times (workload for `ncu --set full`).
    python tools/conv_one.py fprop|dgrad|wgrad n c_in h w c_out k stride pad [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_learning_simulator_b200.ops import conv as C

kind = sys.argv[1]
n, ci, h, w, co, k, st, pad = (int(a) for a in sys.argv[2:10])
iters = int(sys.argv[10]) if len(sys.argv) > 10 else 5
dev = torch.device("cuda")
torch.manual_seed(0)
x = torch.randn(n, ci, h, w, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
wt = (torch.randn(co, ci, k, k, device=dev) / (ci * k * k) ** 0.5).bfloat16().contiguous(memory_format=torch.channels_last)
shape = C.tc_shape(tuple(x.shape), tuple(wt.shape), (st, st), (pad, pad))
oh, ow = (h + 2 * pad - k) // st + 1, (w + 2 * pad - k) // st + 1
dy = torch.randn(n, co, oh, ow, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
for _ in range(iters):
    if kind == "fprop":
        C.tc_fprop(x, wt, shape)
    elif kind == "dgrad":
        C.tc_dgrad(dy, wt, shape)
    else:
        C.tc_wgrad(dy, x, shape, accumulate=True)
torch.cuda.synchronize()
print("ok")
