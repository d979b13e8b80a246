#!/usr/bin/env python
"""One large bf16 GEMM through dtf_gemm_bf16 (for `ncu -k regex:gemm_umma`): wide-MLP forward shape.

    python bench/one_gemm.py [block_n]      block_n = 256 -> CTA-pair kernel, 128 -> single-CTA kernel
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributedtensorflow_b200.ops import kernels as K  # noqa: E402

bn = int(sys.argv[1]) if len(sys.argv) > 1 else 256
M, N, Kd = 4096, 8192, 8192
a = torch.randn(M, Kd, device="cuda").to(torch.bfloat16)
w = torch.randn(Kd, N, device="cuda").to(torch.bfloat16)
out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
for _ in range(3):
    K.gemm(a, w, out, M, N, Kd, b_mn=True, block_n=bn)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
K.gemm(a, w, out, M, N, Kd, b_mn=True, block_n=bn)
e.record()
torch.cuda.synchronize()
ms = s.elapsed_time(e)
print("gemm %dx%dx%d block_n=%d: %.1f us, %.1f TFLOP/s" % (M, N, Kd, bn, ms * 1e3, 2.0 * M * N * Kd / ms / 1e9))
