#!/usr/bin/env python
"""Micro-benchmark of the owner-applies kernels on ONE GPU (everything local): the drain at several
grid caps / pending counts / mirror counts and the publishing push. CUDA-event timed, L2 flushed."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from distributedtensorflow_b200.ops import _lib as L  # noqa: E402
from distributedtensorflow_b200.ops import kernels as K  # noqa: E402


def timeit(fn, reps=20, flush=None):
    ts = []
    for _ in range(reps + 3):
        if flush is not None:
            flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]


def main():
    dev = "cuda"
    n = int(os.environ.get("N", 1638400))
    cols = 1024
    W, depth = 8, 4
    stride = n * 4
    slots = torch.randn(W * depth * n, device=dev) * 1e-3
    flags = torch.zeros(128, dtype=torch.int32, device=dev)
    todo = torch.zeros(32, dtype=torch.int32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    ctrl = torch.zeros(64, dtype=torch.int32, device=dev)
    ctrl.view(torch.float32)[2], ctrl.view(torch.float32)[3] = 0.9, 0.999
    var, m, v = torch.randn(n, device=dev), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    mirrors = [torch.zeros(512 // 4 + n // 2 + 64, dtype=torch.int32, device=dev) for _ in range(W)]
    items = K.make_apply_items([dict(var=var.data_ptr(), slot0=m.data_ptr(), slot1=v.data_ptr(),
                                     shadow=shadow.data_ptr(), grad=slots.data_ptr(), n=n, cols=cols, shadow_ld=cols)])
    mit = K.make_mirror_items([(512, -1, 0)])
    hyper = K.make_hyper(5e-4)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}
    for n_mirror in (0, 2, 8):
        fl64 = flags.view(torch.int64)
        for w in range(W):
            fl64[48 + w] = mirrors[w].data_ptr() if w < n_mirror else 0
        for pending in (1, 2, 8):
            for cap in (0, 32, 64, 148):
                def run():
                    flags[:pending] += 1            # produced
                    K.mailbox_drain(L.OPT_ADAM, items, 1, stride, depth, W, hyper, ctrl.data_ptr(), ctrl.data_ptr(),
                                    flags.data_ptr(), todo.data_ptr(), counter.data_ptr(), None,
                                    mirror_items=mit if n_mirror else None, shard_index=0, max_ctas=cap)
                # time only the drain: bump outside the events
                ts = []
                for _ in range(15):
                    flags[:pending] += 1
                    flush.fill_(1)
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record()
                    K.mailbox_drain(L.OPT_ADAM, items, 1, stride, depth, W, hyper, ctrl.data_ptr(), ctrl.data_ptr(),
                                    flags.data_ptr(), todo.data_ptr(), counter.data_ptr(), None,
                                    mirror_items=mit if n_mirror else None, shard_index=0, max_ctas=cap)
                    e.record()
                    e.synchronize()
                    ts.append(s.elapsed_time(e) * 1e3)
                us = sorted(ts[3:])[len(ts[3:]) // 2]
                byts = n * (24 + 4 * pending + 2 + 2 * n_mirror)
                out["drain_m%d_p%d_c%d" % (n_mirror, pending, cap)] = dict(us=round(us, 2), gbs=round(byts / us / 1e3, 1))
                print("drain mirrors=%d pending=%d cap=%3d: %7.2f us  %7.1f GB/s" % (n_mirror, pending, cap, us, byts / us / 1e3), flush=True)
    # the rmw apply for comparison
    g = torch.randn(n, device=dev) * 1e-3
    it2 = K.make_apply_items([dict(var=var.data_ptr(), slot0=m.data_ptr(), slot1=v.data_ptr(), shadow=shadow.data_ptr(),
                                   grad=g.data_ptr(), n=n, cols=cols, shadow_ld=cols)])
    us = timeit(lambda: K.push_apply(L.OPT_ADAM, it2, 1, hyper, ctrl.data_ptr(), ctrl.data_ptr(), True, 0, None,
                                     counter.data_ptr()), flush=flush)
    print("push_apply (rmw, local): %.2f us  %.1f GB/s" % (us, n * 30 / us / 1e3))
    out["push_apply_rmw"] = dict(us=round(us, 2))
    # publishing push of a few small items
    seq = torch.zeros(1, dtype=torch.int32, device=dev)
    pc = torch.zeros(4, dtype=torch.int32, device=dev)
    minc_dev = torch.zeros(4, dtype=torch.int32, device=dev)
    minc_host = torch.zeros(4, dtype=torch.int32).pin_memory()
    small = torch.randn(64 * 1024, device=dev)
    for label, items_, minc in (("small+dev", [(small.data_ptr(), slots.data_ptr(), 205 * 1024 // 16 * 16, stride)], minc_dev),
                                ("small+host", [(small.data_ptr(), slots.data_ptr(), 205 * 1024 // 16 * 16, stride)], minc_host),
                                ("empty+host", [], minc_host)):
        us = timeit(lambda: K.mailbox_push(items_, depth, 0, [flags.data_ptr()] * 2, True, seq.data_ptr(), pc.data_ptr(),
                                           minc.data_ptr(), my_hdr=mirrors[0].data_ptr()), flush=flush)
        print("publish push %s: %.2f us" % (label, us))
        out["publish_" + label] = dict(us=round(us, 2))
    us = timeit(lambda: K.mailbox_push([(g.data_ptr(), slots.data_ptr(), 12845056, stride)], depth, 0, [flags.data_ptr()], False,
                                       seq.data_ptr(), None, None), flush=flush)
    print("copy push 12.8 MB local: %.2f us" % us)
    zb = torch.zeros(700000, device=dev)
    us = timeit(lambda: K.fill_u32(zb, 0), flush=flush)
    print("fill 2.8 MB: %.2f us" % us)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "micro_mailbox.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
