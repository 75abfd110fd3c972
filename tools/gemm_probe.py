import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_learning_simulator_b200.ops import gemm as G
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conv_check import timeit
dev = torch.device("cuda")
for (m, n, k) in [(1024, 512, 4608), (4096, 256, 2304), (16384, 128, 1152), (65536, 64, 576), (1024, 512, 512), (128, 64, 4608), (128, 64, 64)]:
    x = torch.randn(m, k, device=dev).bfloat16()
    w = torch.randn(n, k, device=dev).bfloat16()
    t = timeit(lambda: G.linear_bf16(x, w, None))
    tc = timeit(lambda: torch.nn.functional.linear(x, w))
    print(f"BN={os.environ.get('DLS_GEMM_BN','auto')} m{m} n{n} k{k}: own {t:.1f} us   cublas {tc:.1f} us", flush=True)
