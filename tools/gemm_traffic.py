"""One launch of each large c3 GEMM shape (after a warm-up launch) for an ncu DRAM-traffic pass:
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:gemm_bf16_kernel python tools/gemm_traffic.py"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from versecrafter_b200 import ops  # noqa: E402

M = 65520
for (N, K) in [(15360, 5120), (5120, 5120), (13824, 5120), (5120, 13824)]:
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / math.sqrt(K)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for _ in range(2):
        ops.gemm(a, w, b, out=out)
    torch.cuda.synchronize()
    del a, w, b, out
