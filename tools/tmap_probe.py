#!/usr/bin/env python3
"""Which activation tensor-map layouts does cuTensorMapEncodeTiled accept?  (debugging aid)"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_learning_simulator_b200.ops import runtime

lib = runtime.require()
buf = torch.zeros(64 * 32 * 32 * 64, dtype=torch.bfloat16, device="cuda")


def probe(name, dims, strides, box, es=None, swizzle=3):
    rank = len(dims)
    es = es or [1] * rank
    d = (ctypes.c_ulonglong * rank)(*dims)
    s = (ctypes.c_ulonglong * max(rank - 1, 1))(*strides)
    b = (ctypes.c_uint * rank)(*box)
    e = (ctypes.c_uint * rank)(*es)
    rc = lib.dls_tensor_map_probe(rank, d, s, b, e, swizzle, buf.data_ptr())
    print(f"{name}: rc={rc} dims={dims} strides={strides} box={box} es={es}", flush=True)


pix = 128
probe("5d s1 as built", [64, 32, 1, 32, 64], [pix, pix * 32, pix * 32, pix * 32 * 32], [64, 32, 1, 4, 1])
probe("5d s1 ph stride = n stride", [64, 32, 1, 32, 64], [pix, pix * 32 * 32 * 64, pix * 32, pix * 32 * 32], [64, 32, 1, 4, 1])
probe("5d s1 dim2 at the end", [64, 32, 32, 64, 1], [pix, pix * 32, pix * 32 * 32, pix * 32 * 32 * 64], [64, 32, 4, 1, 1])
probe("4d s1", [64, 32, 32, 64], [pix, pix * 32, pix * 32 * 32], [64, 32, 4, 1])
probe("5d s2 parity", [128, 16, 2, 16, 64], [pix * 2, pix * 32, pix * 64, pix * 1024], [64, 16, 1, 8, 1])
probe("5d s2 parity, ph after h2", [128, 16, 16, 2, 64], [pix * 2, pix * 64, pix * 32, pix * 1024], [64, 16, 8, 1, 1])
probe("4d elem-stride 2", [64, 32, 32, 64], [pix, pix * 32, pix * 32 * 32], [64, 32, 16, 1], [1, 2, 2, 1])
probe("3d", [64, 1024, 64], [pix, pix * 1024], [64, 128, 1])
probe("5d small box", [64, 32, 1, 32, 64], [pix, pix * 32, pix * 32, pix * 32 * 32], [64, 8, 1, 1, 1])
probe("5d s1 dims2=2", [64, 32, 2, 16, 64], [pix, pix * 32, pix * 64, pix * 32 * 32], [64, 32, 1, 4, 1])
