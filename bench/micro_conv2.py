#!/usr/bin/env python
"""conv2 forward / dgrad: halo-tile kernels against the tap-by-tap ones (CUDA events, L2 flushed)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from distributedtensorflow_b200.ops import kernels as K

def timeit(fn, flush, reps=30):
    ts = []
    for _ in range(reps + 5):
        flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts = sorted(ts[5:]); return ts[len(ts) // 2]

B = 128
dev = "cuda"
pool1p = torch.zeros(B, 18, 18, 32, dtype=torch.bfloat16, device=dev)
pool1p[:, 2:16, 2:16] = torch.randn(B, 14, 14, 32, device=dev).to(torch.bfloat16)
w2 = (torch.randn(800, 64, device=dev) * 0.05).to(torch.bfloat16)
b2 = torch.randn(64, device=dev) * 0.1
pool2 = torch.zeros(B, 7, 7, 64, dtype=torch.bfloat16, device=dev)
amax2 = torch.zeros(B, 7, 7, 64, dtype=torch.uint8, device=dev)
dconv2p = torch.zeros(B, 18, 18, 64, dtype=torch.bfloat16, device=dev)
dconv2p[:, 2:16, 2:16] = torch.randn(B, 14, 14, 64, device=dev).to(torch.bfloat16)
dpool1 = torch.zeros(B, 14, 14, 32, dtype=torch.bfloat16, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
empty = timeit(lambda: K.fill_u32(torch.zeros(64, device=dev), 0), flush)
for halo in (False, True):
    f = timeit(lambda: K.conv2_fwd(pool1p, w2.data_ptr(), b2, pool2, amax2, halo=halo), flush)
    d = timeit(lambda: K.conv2_dgrad(dconv2p, w2.data_ptr(), dpool1, halo=halo), flush)
    print("halo=%s  fwd %.2f us  dgrad %.2f us   (tiny-kernel floor of this harness %.2f us)" % (halo, f, d, empty))
