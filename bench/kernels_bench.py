#!/usr/bin/env python
"""Per-kernel micro-benchmarks with roofline fractions against MEASURED_PEAKS.json
(hbm_gbs, bf16_tflops; fallback 6650 GB/s / 1590 TFLOP/s).

    python bench/kernels_bench.py [--out gpurun_out/kernels_bench.json]

* tcgen05 GEMM at the wide-MLP shapes (BASELINE config 4: batch 4096, hidden 8192) for the three
  layouts (fwd / dX / dW), next to torch.matmul (cuBLAS) on the same shapes;
* push+apply (Adam / SGD / atomic SGD) and pull_bulk on a local shard: achieved GB/s vs the HBM
  copy peak (the NVLink variants are measured by bench/pushpull_sweep.py);
* the CNN's conv kernels.
Timing: CUDA events, 3 warm-ups, L2 flushed (256 MiB write) between timed iterations.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from distributedtensorflow_b200.ops import _lib as L  # noqa: E402
from distributedtensorflow_b200.ops import kernels as K  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


_flush = None


def timeit(fn, iters=10, warmup=3):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for i in range(iters):
        _flush.fill_(i)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]  # median, microseconds


def bench_gemm(results, hbm, tf):
    bf = torch.bfloat16
    shapes = [("wide fwd  Y=XW", 4096, 8192, 8192, False, True),
              ("wide dX   dY W^T", 4096, 8192, 8192, False, False),
              ("wide dW   X^T dY", 8192, 8192, 4096, True, True),
              ("wide fwd0 784->8192", 4096, 8192, 784, False, True),
              ("cnn dense1 fwd", 128, 1024, 3136, False, True),
              ("cnn dense1 dW", 3136, 1024, 128, True, True)]
    for name, M, N, Kd, a_mn, b_mn in shapes:
        A = torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda").to(bf)
        B = torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda").to(bf)
        out = torch.empty(M, N, dtype=bf, device="cuda")
        for bn in ((128, 256) if M >= 1024 else (128,)):
            us = timeit(lambda: K.gemm(A, B, out, M, N, Kd, a_mn=a_mn, b_mn=b_mn, block_n=bn))
            a_t = A.t() if a_mn else A
            b_t = B if b_mn else B.t()
            us_ref = timeit(lambda: torch.matmul(a_t, b_t))
            flops = 2.0 * M * N * Kd
            byts = 2.0 * (M * Kd + N * Kd + M * N)
            roof_us = max(flops / (tf * 1e6), byts / (hbm * 1e3))
            results.append(dict(kernel="gemm_umma", case=name, M=M, N=N, K=Kd, block_n=bn, us=round(us, 2),
                                tflops=round(flops / us / 1e6, 1), cublas_us=round(us_ref, 2),
                                cublas_tflops=round(flops / us_ref / 1e6, 1),
                                frac_of_roofline=round(roof_us / us, 3), vs_cublas=round(us_ref / us, 3)))
            print(results[-1], flush=True)


def bench_apply_pull(results, hbm):
    n = 64 << 20   # 64 Mi parameters = 256 MiB fp32
    var = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    g = torch.randn(n, device="cuda")
    cols = 8192
    shadow = torch.zeros(n // cols, cols, dtype=torch.bfloat16, device="cuda")
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")
    ctrl.view(torch.float32)[2], ctrl.view(torch.float32)[3] = 0.9, 0.999
    step = torch.zeros(1, dtype=torch.int64, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    items = K.make_apply_items([dict(var=var.data_ptr(), slot0=m.data_ptr(), slot1=v.data_ptr(),
                                     shadow=shadow.data_ptr(), grad=g.data_ptr(), n=n, cols=cols, shadow_ld=cols)])
    for name, opt, atomic, byts in (("adam", L.OPT_ADAM, False, n * (4 + 12 + 12 + 2)),
                                    ("sgd", L.OPT_SGD, False, n * (4 + 4 + 4 + 2)),
                                    ("sgd red.add", L.OPT_SGD, True, n * (4 + 4))):
        hyper = K.make_hyper(1e-3, atomic_sgd=atomic)
        us = timeit(lambda: K.push_apply(opt, items, 1, hyper, ctrl.data_ptr(), ctrl.data_ptr(), True, 0,
                                         step.data_ptr(), counter.data_ptr()), iters=5)
        results.append(dict(kernel="push_apply", case=name + " local shard", elements=n, us=round(us, 1),
                            gbs=round(byts / us / 1e3, 1), frac_of_hbm=round(byts / us / 1e3 / hbm, 3)))
        print(results[-1], flush=True)
    dst = torch.empty_like(var)
    us = timeit(lambda: K.pull_bulk([(var.data_ptr(), dst.data_ptr(), n * 4)]), iters=5)
    results.append(dict(kernel="pull_bulk", case="local shard 256 MiB", us=round(us, 1),
                        gbs=round(2 * n * 4 / us / 1e3, 1), frac_of_hbm=round(2 * n * 4 / us / 1e3 / hbm, 3)))
    print(results[-1], flush=True)
    us = timeit(lambda: dst.copy_(var), iters=5)
    results.append(dict(kernel="torch copy_ (reference)", case="256 MiB", us=round(us, 1),
                        gbs=round(2 * n * 4 / us / 1e3, 1)))
    print(results[-1], flush=True)


def bench_conv(results, tf):
    B = 128
    bf = torch.bfloat16
    pool1p = torch.zeros(B, 18, 18, 32, dtype=bf, device="cuda")
    pool1p[:, 2:16, 2:16] = torch.rand(B, 14, 14, 32, device="cuda").to(bf)
    w2 = (torch.randn(800, 64, device="cuda") * 0.05).to(bf)
    b2 = torch.zeros(64, device="cuda")
    pool2 = torch.zeros(B, 7, 7, 64, dtype=bf, device="cuda")
    amax2 = torch.zeros(B, 7, 7, 64, dtype=torch.uint8, device="cuda")
    dconv2p = torch.zeros(B, 18, 18, 64, dtype=bf, device="cuda")
    dconv2p[:, 2:16, 2:16] = torch.randn(B, 14, 14, 64, device="cuda").to(bf)
    dw = torch.zeros(800, 64, device="cuda")
    dpool1 = torch.zeros(B, 14, 14, 32, dtype=bf, device="cuda")
    flops = 2.0 * B * 196 * 800 * 64
    for name, fn in (("conv2_fwd", lambda: K.conv2_fwd(pool1p, w2.data_ptr(), b2, pool2, amax2)),
                     ("conv2_wgrad", lambda: K.conv2_wgrad(pool1p, dconv2p, dw)),
                     ("conv2_dgrad", lambda: K.conv2_dgrad(dconv2p, w2.data_ptr(), dpool1))):
        us = timeit(fn)
        results.append(dict(kernel=name, case="B=128", us=round(us, 2), tflops=round(flops / us / 1e6, 1),
                            frac_of_bf16_peak=round(flops / us / 1e6 / tf, 4)))
        print(results[-1], flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "kernels_bench.json"))
    args = ap.parse_args()
    hbm, tf, src = peaks()
    results = [dict(peaks=dict(hbm_gbs=hbm, bf16_tflops=tf, source=src))]
    bench_gemm(results, hbm, tf)
    bench_apply_pull(results, hbm)
    bench_conv(results, tf)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(results, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
