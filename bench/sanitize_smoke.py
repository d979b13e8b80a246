#!/usr/bin/env python
"""Small driver for ``compute-sanitizer``: one CNN training step through the engine
(every conv / GEMM / loss / push+apply / pull kernel), one split-K GEMM with the in-kernel fix-up,
and one CTA-pair GEMM. Sizes are tiny so the 10-50x sanitizer slowdown stays in seconds.

    compute-sanitizer --tool memcheck  python bench/sanitize_smoke.py
    compute-sanitizer --tool racecheck python bench/sanitize_smoke.py     (shared-memory hazards)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import __graft_entry__ as entry  # noqa: E402
from distributedtensorflow_b200.ops import kernels as K  # noqa: E402


def main() -> None:
    entry.smoke()
    bf = torch.bfloat16
    torch.manual_seed(0)
    # split-K with fix-up (fp32 workspace + self-resetting tile counters)
    M, N, Kd = 128, 256, 1024
    a = torch.randn(M, Kd, device="cuda").to(bf)
    w = torch.randn(Kd, N, device="cuda").to(bf)
    out = torch.zeros(M, N, dtype=bf, device="cuda")
    ws = torch.zeros(M, N, device="cuda")
    cnt = torch.zeros(16, dtype=torch.int32, device="cuda")
    K.gemm(a, w, out, M, N, Kd, b_mn=True, relu=True, splits=4, block_n=64, ws=ws, counters=cnt)
    ref = torch.relu(a.float() @ w.float())
    torch.cuda.synchronize()
    assert (out.float() - ref).norm() / ref.norm() < 1e-2
    # CTA-pair kernel, edge tiles in M, N and K
    M, N, Kd = 300, 520, 200
    a = torch.randn(M, Kd, device="cuda").to(bf)
    b = torch.randn(N, Kd, device="cuda").to(bf)
    out = torch.zeros(M, N, device="cuda")
    K.gemm(a, b, out, M, N, Kd, block_n=256)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    assert (out - ref).norm() / ref.norm() < 2e-3
    print("sanitize_smoke: ok")


if __name__ == "__main__":
    main()
