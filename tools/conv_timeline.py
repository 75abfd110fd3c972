#!/usr/bin/env python3
"""Per-CTA phase stamps (clock64) of the halo-patch convolution kernel.  Needs a library built with
``DLS_NVCC_EXTRA=-DDLS_CONV_TIMELINE``:

    DLS_NVCC_EXTRA=-DDLS_CONV_TIMELINE python -m distributed_learning_simulator_b200.build --force
    python tools/conv_timeline.py fprop 64 64 32 32 64
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_learning_simulator_b200.ops import conv as C
from distributed_learning_simulator_b200.ops import runtime

kind = sys.argv[1]
n, ci, h, w, co = (int(a) for a in sys.argv[2:7])
dev = torch.device("cuda")
x = torch.randn(n, ci, h, w, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
wt = (torch.randn(co, ci, 3, 3, device=dev) / (ci * 9) ** 0.5).bfloat16().contiguous(memory_format=torch.channels_last)
shape = C.tc_shape(tuple(x.shape), tuple(wt.shape), (1, 1), (1, 1))
dy = torch.randn(n, co, h, w, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
lib = ctypes.CDLL(runtime.lib_path())
lib.dls_conv_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
run = (lambda: C.tc_fprop(x, wt, shape)) if kind == "fprop" else (lambda: C.tc_dgrad(dy, wt, shape))
for _ in range(5):
    run()
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (148 * 16))()
names = ["entry", "setup", "pdl_wait", "first_A", "t0_acc", "t0_done", "t1_acc", "t1_done", "t2_acc", "t2_done",
         "t3_acc", "t3_done", "t0_ld0", "t0_st0", "end"]
assert lib.dls_conv_timeline(buf, 0) == 0
for cta in (0, 1, 73, 147):
    row = [buf[cta * 16 + i] for i in range(15)]
    base = row[0]
    print(f"cta {cta}: " + "  ".join(f"{nm}={row[i] - base}" for i, nm in enumerate(names) if row[i]))
