#!/usr/bin/env python
"""BASELINE config 5: parameter push / pull bandwidth sweep, 4 KB - 256 MB (1 GB with --max-mb
1024), our in-kernel NVLink paths vs NCCL on the same pairs of GPUs.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench/pushpull_sweep.py [--max-mb 256] [--out gpurun_out/sweep.json]

Per size S and rank r (peer = (r+1) % N):
  pull        dtf_pull_bulk: S bytes peer arena -> local (16-byte peer loads)
  push_sgd    dtf_push_apply SGD with red.global.add.f32 into the peer's variable (write-only)
  push_adam   dtf_push_apply Adam: remote read-modify-write of var/m/v + bf16 serving copy
  pull_all    S/N bytes from EACH of the N shards (the flat-sliced layout's pull)
  mbox_push   dtf_mailbox_push: S bytes of gradient into my mailbox on the peer (posted 16-byte stores) -
              the DEFAULT multi-GPU push path; mbox_publish_us = the same with the publishing tail
              (one system-scope fence, st.release of the sequence number)
  drain       dtf_mailbox_drain_ex on MY shard: Adam from local HBM on S/4 parameters with one pending
              gradient, no mirror (drain_us) and with the peer's mirror registered (drain_mirror_us):
              mirror_extra_us = what the posted bf16 stores into the peer's mirror add to the drain
  gloo_p2p    ring send/recv of S bytes between host tensors over gloo/TCP (what the CPU transport and
              the reference's gRPC data plane resemble), wall clock, sizes <= 64 MB
  nccl_p2p    ring send/recv of S bytes (batch_isend_irecv)
  nccl_ag     all_gather producing S bytes per rank       (NCCL "pull" of a sharded variable)
  nccl_rs     reduce_scatter consuming S bytes per rank   (NCCL "push" of a sharded gradient)
GB/s = payload bytes that cross NVLink per rank / device time (CUDA events, max over ranks,
3 warm-ups, 5 timed, L2 flushed between iterations); the pulled bytes are compared byte-exactly
with what NCCL delivered. Roofline = 770 GB/s measured peer copy (900 GB/s nominal).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NVLINK_MEASURED_GBS = 770.0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from distributedtensorflow_b200.ops import _lib as L
    from distributedtensorflow_b200.ops import kernels as K

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = L.lib()
    max_bytes = args.max_mb << 20
    n_max = max_bytes // 4

    # arena: [ctrl | var | m | v | shadow(bf16)] exported over CUDA IPC to every rank
    ctrl_b = 4096
    # off_var is only ever pulled (so it can be compared with NCCL); pushes target off_pv/m/v/sh
    off_var, off_pv = ctrl_b, ctrl_b + max_bytes
    off_m, off_v = ctrl_b + 2 * max_bytes, ctrl_b + 3 * max_bytes
    off_sh = ctrl_b + 4 * max_bytes
    off_mb = off_sh + max_bytes // 2 + 4096            # my peers' mailbox inside MY arena (one buffer)
    off_flags = off_mb + max_bytes                     # dtf_mbox_flags
    off_mir = off_flags + 4096                         # MY mirror (header + bf16 copy), written by my peer's drain
    arena_bytes = off_mir + 512 + max_bytes // 2 + 4096
    ptr = C.c_void_p()
    L.check(lib.dtf_arena_alloc(C.byref(ptr), arena_bytes), "arena_alloc")
    handle = C.create_string_buffer(L.IPC_HANDLE_BYTES)
    L.check(lib.dtf_ipc_export(ptr, handle), "ipc_export")
    handles = [None] * world
    dist.all_gather_object(handles, bytes(handle.raw))
    bases = []
    for r in range(world):
        if r == rank:
            bases.append(int(ptr.value))
        else:
            p = C.c_void_p()
            L.check(lib.dtf_ipc_import(handles[r], C.byref(p)), "ipc_import")
            bases.append(int(p.value))
    # fill my variable with a rank-specific pattern through a torch view of the arena
    pattern = (torch.arange(n_max, device="cuda", dtype=torch.float32) % 1024) + rank * 1000.0
    L.check(lib.dtf_memcpy_h2d(bases[rank] + off_var, pattern.cpu().numpy().ctypes.data_as(C.c_void_p),
                               max_bytes, None), "h2d")
    torch.cuda.synchronize()
    dist.barrier()

    local = torch.zeros(n_max, device="cuda")
    grad = torch.randn(n_max, device="cuda")
    recv = torch.zeros(n_max, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    peer = (rank + 1) % world
    prev = (rank - 1) % world
    gloo = dist.new_group(backend="gloo")
    host_send = torch.empty(min(max_bytes, 64 << 20), dtype=torch.uint8)
    host_recv = torch.empty_like(host_send)
    seq = torch.zeros(4, dtype=torch.int32, device="cuda")
    pcount = torch.zeros(4, dtype=torch.int32, device="cuda")
    minc = torch.zeros(4, dtype=torch.int32, device="cuda")
    todo = torch.zeros(32, dtype=torch.int32, device="cuda")
    dcount = torch.zeros(4, dtype=torch.int32, device="cuda")
    one_u32 = (C.c_uint * 1)(1)

    def timed(fn, iters=5, warmup=3):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        tot = []
        for i in range(iters):
            flush.fill_(i)
            torch.cuda.synchronize()
            dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            tot.append(s.elapsed_time(e) * 1e3)
        t = torch.tensor([sorted(tot)[len(tot) // 2]], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())     # microseconds, max over ranks of the per-rank median

    results = []
    size = 4096
    while size <= max_bytes:
        n = size // 4
        row = dict(bytes=size, n_gpus=world)

        def pull():
            K.pull_bulk([(bases[peer] + off_var, local.data_ptr(), size)])
        us = timed(pull)
        row["pull_gbs"] = round(size / us / 1e3, 2)

        # byte-exact check against NCCL p2p of the same bytes
        mine = pattern[:n].contiguous()
        ops = [dist.P2POp(dist.isend, mine, prev), dist.P2POp(dist.irecv, recv[:n], peer)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        torch.cuda.synchronize()
        row["pull_equals_nccl"] = bool(torch.equal(local[:n], recv[:n]))

        def nccl_p2p():
            o = [dist.P2POp(dist.isend, mine, prev), dist.P2POp(dist.irecv, recv[:n], peer)]
            for w in dist.batch_isend_irecv(o):
                w.wait()
        us = timed(nccl_p2p)
        row["nccl_p2p_gbs"] = round(size / us / 1e3, 2)

        hyper = K.make_hyper(1e-3, atomic_sgd=True)
        items_sgd = K.make_apply_items([dict(var=bases[peer] + off_pv, grad=grad.data_ptr(), n=n, cols=1)])

        def push_sgd():
            K.push_apply(L.OPT_SGD, items_sgd, 1, hyper, None, bases[peer], False, rank, None, counter.data_ptr())
        us = timed(push_sgd)
        row["push_sgd_red_gbs"] = round(size / us / 1e3, 2)

        hyper_a = K.make_hyper(1e-3)
        cols = 1024 if n % 1024 == 0 else 1
        items_adam = K.make_apply_items([dict(var=bases[peer] + off_pv, slot0=bases[peer] + off_m,
                                              slot1=bases[peer] + off_v, shadow=bases[peer] + off_sh,
                                              grad=grad.data_ptr(), n=n, cols=cols, shadow_ld=cols)])

        def push_adam():
            K.push_apply(L.OPT_ADAM, items_adam, 1, hyper_a, bases[peer], bases[peer], False, rank, None,
                         counter.data_ptr())
        us = timed(push_adam)
        row["push_adam_rmw_us"] = round(us, 2)
        row["push_adam_link_gbs_in"] = round(12.0 * n / us / 1e3, 2)    # remote reads
        row["push_adam_link_gbs_out"] = round(14.0 * n / us / 1e3, 2)   # remote writes

        # -- the default push path: posted stores into my mailbox on the peer, then the publishing tail
        def mbox_push():
            K.mailbox_push([(grad.data_ptr(), bases[peer] + off_mb, size, max_bytes)], 1, 0,
                           [bases[peer] + off_flags], False, seq.data_ptr(), None, None)
        us = timed(mbox_push)
        row["mbox_push_gbs"] = round(size / us / 1e3, 2)

        def mbox_publish():
            K.mailbox_push([(grad.data_ptr(), bases[peer] + off_mb, size, max_bytes)], 1, 0,
                           [bases[peer] + off_flags], True, seq.data_ptr(), pcount.data_ptr(), minc.data_ptr())
        row["mbox_push_publish_us"] = round(timed(mbox_publish), 2)
        row["mbox_push_frac_of_900"] = round(row["mbox_push_gbs"] / 900.0, 3)

        # -- the owner side: drain my shard (Adam, one pending gradient), without / with the peer's mirror
        d_items = K.make_apply_items([dict(var=bases[rank] + off_pv, slot0=bases[rank] + off_m, slot1=bases[rank] + off_v,
                                           shadow=bases[rank] + off_sh, grad=bases[rank] + off_mb, n=n, cols=cols,
                                           shadow_ld=cols)])
        d_mirror = K.make_mirror_items([(512, -1, 0)])
        flags_ptr = bases[rank] + off_flags

        def set_mirror(addr):
            buf = (C.c_ulonglong * 1)(addr)
            L.check(lib.dtf_memcpy_h2d(flags_ptr + L.MBOX_OFF_MIRROR_BASE, buf, 8, None), "h2d")
            L.check(lib.dtf_stream_sync(None), "sync")

        def drain(with_mirror):
            def f():
                # one more gradient "arrives": produced[0] += 1 (host write, outside nothing: part of the call)
                L.check(lib.dtf_memcpy_h2d(flags_ptr + L.MBOX_OFF_PRODUCED, prod_buf, 4, None), "h2d")
                K.mailbox_drain(L.OPT_ADAM, d_items, 1, max_bytes, 1, 1, hyper_a, bases[rank], None, flags_ptr,
                                todo.data_ptr(), dcount.data_ptr(), None,
                                mirror_items=d_mirror if with_mirror else None, shard_index=0)
                prod_buf[0] += 1
            return f
        zero512 = (C.c_ubyte * 512)()
        L.check(lib.dtf_memcpy_h2d(flags_ptr, zero512, 512, None), "h2d")
        L.check(lib.dtf_stream_sync(None), "sync")
        prod_buf = (C.c_uint * 1)(1)
        row["drain_us"] = round(timed(drain(False)), 2)
        if world > 1:
            set_mirror(bases[peer] + off_mir)
            row["drain_mirror_us"] = round(timed(drain(True)), 2)
            set_mirror(0)
            # the mirror stores overlap the drain's own HBM traffic: report what they add, and the rate
            # at which the drain as a whole feeds the link
            row["mirror_bytes"] = 2 * n
            row["mirror_extra_us"] = round(row["drain_mirror_us"] - row["drain_us"], 2)
            row["mirror_link_gbs_during_drain"] = round(2.0 * n / row["drain_mirror_us"] / 1e3, 2)

        if size <= host_send.numel():
            import time as _t
            hs, hr = host_send[:size], host_recv[:size]

            def gloo_once():
                reqs = [dist.isend(hs, prev, group=gloo), dist.irecv(hr, peer, group=gloo)]
                for q in reqs:
                    q.wait()
            for _ in range(2):
                gloo_once()
            dist.barrier()
            t0 = _t.perf_counter()
            for _ in range(3):
                gloo_once()
            gus = (_t.perf_counter() - t0) / 3 * 1e6
            tg = torch.tensor([gus], device="cuda", dtype=torch.float64)
            dist.all_reduce(tg, op=dist.ReduceOp.MAX)
            row["gloo_p2p_gbs"] = round(size / float(tg.item()) / 1e3, 3)

        if world > 1 and size >= 16 * world:
            per = (size // world) // 16 * 16

            def pull_all():
                # ring order starting at my neighbour: at any moment the N ranks read N different peers
                K.pull_bulk([(bases[s] + off_var, local.data_ptr() + s * per, per)
                             for s in [(rank + k) % world for k in range(1, world)]])
            us = timed(pull_all)
            row["pull_all_shards_gbs"] = round(per * (world - 1) / us / 1e3, 2)
            ag_in = pattern[:per // 4].contiguous()
            ag_out = torch.empty(per // 4 * world, device="cuda")

            def nccl_ag():
                dist.all_gather_into_tensor(ag_out, ag_in)
            us = timed(nccl_ag)
            row["nccl_all_gather_gbs"] = round(per * (world - 1) / us / 1e3, 2)
            rs_in = grad[:per // 4 * world].contiguous()
            rs_out = torch.empty(per // 4, device="cuda")

            def nccl_rs():
                dist.reduce_scatter_tensor(rs_out, rs_in)
            us = timed(nccl_rs)
            row["nccl_reduce_scatter_gbs"] = round(per * (world - 1) / us / 1e3, 2)
        row["pull_frac_of_900"] = round(row["pull_gbs"] / 900.0, 3)
        row["pull_frac_of_nvlink"] = round(row["pull_gbs"] / NVLINK_MEASURED_GBS, 3)
        row["push_sgd_frac_of_nvlink"] = round(row["push_sgd_red_gbs"] / NVLINK_MEASURED_GBS, 3)
        results.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        size *= 4
    if rank == 0:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump(results, open(args.out, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
