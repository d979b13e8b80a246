"""Host-side layout of the B200 data plane (no GPU needed): shard arenas, mailboxes, worker mirrors,
and the byte offsets the Python side shares with the C structs of include/dtf_b200.h."""
import ctypes as C
import re
import os

import numpy as np

from distributedtensorflow_b200.ops import _lib as L
from distributedtensorflow_b200.parallel import nvl
from distributedtensorflow_b200.parallel.layout import VarLayout, VarSlice
from distributedtensorflow_b200.train.optimizers import AdamOptimizer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _layout(num_ps):
    """conv kernel (small 4-D), a flat-sliced dense kernel, its bias, a 10-column head, global_step."""
    def whole(name, shape, shard, idx, dtype="float32", trainable=True):
        n = int(np.prod(shape)) if shape else 1
        return VarLayout(name, tuple(shape), dtype, trainable, [VarSlice(shard, 0, shape[0] if shape else 1, 0, n)], idx)
    rows, cols = 512, 256
    per = rows // num_ps
    dense = VarLayout("dense/kernel", (rows, cols), "float32", True,
                      [VarSlice(s, s * per, (s + 1) * per, s * per * cols, per * cols) for s in range(num_ps)], 1)
    lay = {"conv/kernel": whole("conv/kernel", (5, 5, 1, 32), 0, 0),
           "dense/kernel": dense,
           "dense/bias": whole("dense/bias", (256,), 1 % num_ps, 2),
           "head/kernel": whole("head/kernel", (256, 10), 0, 3),
           "global_step": whole("global_step", (), 0, 4, dtype="int64", trainable=False)}
    return lay


def test_c_struct_sizes_and_offsets_match_the_header():
    text = open(os.path.join(ROOT, "include", "dtf_b200.h")).read()

    def body(struct):
        a = text.index("typedef struct %s {" % struct)
        return text[a:text.index("} %s;" % struct, a)]
    hdr = body("dtf_mbox_flags")
    assert C.sizeof(L.DrainMirrorItem) == 24 and C.sizeof(L.MboxPushItem) == 40 and C.sizeof(L.PsCtrl) == 256
    # offsets quoted in the struct comments are the ones Python uses
    for name, off in (("lock", L.MBOX_OFF_LOCK), ("paused", L.MBOX_OFF_PAUSED), ("drains", L.MBOX_OFF_DRAINS),
                      ("applied_total", L.MBOX_OFF_APPLIED), ("tag0[DTF_MAX_WORKERS]", L.MBOX_OFF_TAG0),
                      ("tag1[DTF_MAX_WORKERS]", L.MBOX_OFF_TAG1), ("consumed[DTF_MAX_WORKERS]", L.MBOX_OFF_CONSUMED),
                      ("done[DTF_MAX_WORKERS]", L.MBOX_OFF_DONE),
                      ("mirror_base[DTF_MAX_WORKERS]", L.MBOX_OFF_MIRROR_BASE)):
        m = re.search(r"unsigned (?:int|long long) %s;\s*/\* \+(\d+)" % re.escape(name), hdr)
        assert m and int(m.group(1)) == off, (name, off, m and m.group(1))
    hdr = body("dtf_mirror_hdr")
    for name, off in (("global_step", L.MIRROR_OFF_GLOBAL_STEP), ("epoch", L.MIRROR_OFF_EPOCH),
                      ("abort", L.MIRROR_OFF_ABORT), ("version[DTF_MIRROR_MAX_SHARDS]", L.MIRROR_OFF_VERSION),
                      ("consumed[DTF_MIRROR_MAX_SHARDS]", L.MIRROR_OFF_CONSUMED)):
        m = re.search(r"unsigned (?:int|long long) %s;\s*/\* \+(\d+)" % re.escape(name), hdr)
        assert m and int(m.group(1)) == off, (name, off, m and m.group(1))
    # produced 64 + tag0 64 + consumed 64 + tag1 64 + done 64 + 4 words + 12 pad + 16 x 8 = 512
    assert 5 * 64 + 16 + 48 + 128 == L.MBOX_FLAGS_BYTES
    assert L.MIRROR_OFF_CONSUMED + 4 * L.MIRROR_MAX_SHARDS <= L.MIRROR_HDR_BYTES


def test_arena_layout_with_mailboxes_is_aligned_and_disjoint():
    spec = AdamOptimizer(0.001).spec()
    lay = _layout(2)
    names = [n for n, v in lay.items() if v.trainable]
    al = nvl.build_arena_layout(lay, 2, spec, names, sync_workers=3, mailbox_depth=4, mirrors=False)
    assert al.step_var == "global_step" and al.mailbox_depth == 4 and not al.mirror_f32
    for sh in range(2):
        spans = []
        for name, regs in al.regions.items():
            for r in regs:
                if r.shard != sh:
                    continue
                spans.append((r.var_off, r.var_off + r.sl.n * 4))
                for so in r.slot_off:
                    spans.append((so, so + r.sl.n * 4))
                if r.shadow_off is not None:
                    spans.append((r.shadow_off, r.shadow_off + r.shadow_bytes))
        spans.append((al.mbox_flags_off[sh], al.mbox_flags_off[sh] + L.MBOX_FLAGS_BYTES))
        slots_end = al.mbox_slot_base[sh] + 3 * 4 * al.mailbox_stride[sh]
        spans.append((al.mbox_slot_base[sh], slots_end))
        spans.sort()
        assert spans[0][0] >= L.CTRL_BYTES
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 <= b0, (sh, (a0, a1), (b0, b1))
        assert all(a % 256 == 0 for a, _ in spans)
        assert al.shard_bytes[sh] >= slots_end
    # every trainable slice has its place inside a mailbox slot of its owner
    for name in names:
        for r in al.regions[name]:
            off = al.mailbox_off[(name, r.shard)] - al.mbox_slot_base[r.shard]
            assert 0 <= off and off + r.sl.n * 4 <= al.mailbox_stride[r.shard]


def test_worker_mirror_layout_holds_the_forms_the_kernels_read():
    spec = AdamOptimizer(0.001).spec()
    lay = _layout(4)
    names = [n for n, v in lay.items() if v.trainable]
    al = nvl.build_arena_layout(lay, 4, spec, names, sync_workers=2, mailbox_depth=2, mirrors=False)
    ml = nvl.build_mirror_layout(lay, al.regions)
    # matrices: bf16 row-padded copy; vectors and small tensors: fp32 copy; a small matrix gets both
    assert set(ml.bf16) == {"conv/kernel", "dense/kernel", "head/kernel"}
    assert set(ml.f32) == {"conv/kernel", "dense/bias", "head/kernel"}
    assert "global_step" not in ml.f32 and "global_step" not in ml.bf16
    spans = [(L.MIRROR_OFF_GLOBAL_STEP, L.MIRROR_HDR_BYTES)]
    for n, o in ml.f32.items():
        spans.append((o, o + lay[n].num_elements * 4))
    for n, o in ml.bf16.items():
        ld = al.regions[n][0].shadow_ld
        assert ld % 8 == 0 and ld >= lay[n].cols
        spans.append((o, o + lay[n].rows * ld * 2))
    spans.sort()
    for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
        assert a1 <= b0 and b0 % 256 == 0
    assert ml.nbytes >= spans[-1][1] and ml.nbytes % 4096 == 0
    # the 10-column head is padded to 16 columns in its bf16 form
    assert al.regions["head/kernel"][0].shadow_ld == 16


def test_every_ctypes_signature_matches_the_header_prototype():
    """ABI drift guard: each function bound in ops/_lib.py exists in include/dtf_b200.h with the same
    number of parameters (the bindings are hand-written; a missed argument shifts every pointer after it)."""
    import inspect
    text = open(os.path.join(ROOT, "include", "dtf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)          # drop comments (they contain commas / parens)
    protos = {}
    for m in re.finditer(r"\bint\s+(dtf_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
    src = inspect.getsource(L._declare)
    bound = re.findall(r'"(dtf_\w+)":\s*\[', src)
    assert len(bound) > 40
    ns = {"vp": C.c_void_p, "i": C.c_int, "f": C.c_float, "ll": C.c_longlong, "C": C}
    ns.update({k: getattr(L, k) for k in dir(L)})
    sigs_src = src[src.index("sigs = {"):src.index("for name, args in sigs.items()")]
    exec("import ctypes as C\n" + inspect.cleandoc(sigs_src), ns)
    for name in bound:
        assert name in protos, "%s is bound but not declared in include/dtf_b200.h" % name
        assert len(ns["sigs"][name]) == protos[name], (name, len(ns["sigs"][name]), protos[name])


def test_host_library_bindings_match_their_header():
    import inspect
    from distributedtensorflow_b200.utils import hostlib
    text = open(os.path.join(ROOT, "include", "dtf_b200_io.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(dtf_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
    src = inspect.getsource(hostlib._declare)
    for name, arglist in re.findall(r'\("(dtf_\w+)",\s*\w+,\s*\[(.*?)\]\),?\n', src, flags=re.S):
        n = 0 if not arglist.strip() else len(re.sub(r"\([^()]*\)", "", arglist).split(","))
        assert name in protos, "%s is bound but not declared in include/dtf_b200_io.h" % name
        assert n == protos[name], (name, n, protos[name])
