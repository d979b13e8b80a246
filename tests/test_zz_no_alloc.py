"""``dtf_push_apply_static``: the push+apply entry point for callers that cannot allocate device
memory — a PS task whose HBM is fully committed to its shard arena, a host thread that must stay
out of ``cudaMalloc`` while kernels are in flight, or a foreign runtime that only has the C ABI.
Its CTA-election scratch lives in a ``__device__`` array inside the library instead of being
supplied by the caller.

The test drives BOTH entry points through the C ABI (ctypes on the shared object, no Python
wrapper in between), on identical copies of the same optimizer state, twice in a row, and
requires the same results; the second call proves the shared scratch is re-armed between calls.
The file sorts last on purpose: with ``pytest -x`` every other test has already run when this
one is reached. It passes both alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(lib):
    from distributedtensorflow_b200.ops import _lib as L
    vp, i = C.c_void_p, C.c_int
    lib.dtf_push_apply_static.argtypes = [i, C.POINTER(L.ApplyItem), i, C.POINTER(L.ApplyHyper), vp, vp,
                                          i, i, vp, vp]
    lib.dtf_push_apply_static.restype = i
    return L


def _state(shapes, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    out = []
    for shp in shapes:
        n = 1
        for d in shp:
            n *= d
        cols = shp[-1]
        ld = (cols + 7) // 8 * 8
        out.append(dict(n=n, cols=cols, ld=ld,
                        var=torch.randn(n, device="cuda", generator=g),
                        m=torch.rand(n, device="cuda", generator=g) * 0.1,
                        v=torch.rand(n, device="cuda", generator=g) * 0.01,
                        shadow=torch.zeros(n // cols, ld, dtype=torch.bfloat16, device="cuda")))
    return out


def _clone(state):
    return [{k: (t.clone() if isinstance(t, torch.Tensor) else t) for k, t in s.items()} for s in state]


def _items(L, state, grads):
    arr = (L.ApplyItem * len(state))()
    for k, (s, g) in enumerate(zip(state, grads)):
        arr[k].var, arr[k].slot0, arr[k].slot1 = s["var"].data_ptr(), s["m"].data_ptr(), s["v"].data_ptr()
        arr[k].shadow, arr[k].grad = s["shadow"].data_ptr(), g.data_ptr()
        arr[k].n, arr[k].first, arr[k].cols, arr[k].shadow_ld = s["n"], 0, s["cols"], s["ld"]
    return arr


def _ctrl():
    c = torch.zeros(64, dtype=torch.int32, device="cuda")       # one 256-byte dtf_ps_ctrl
    c.view(torch.float32)[2], c.view(torch.float32)[3] = 0.9, 0.999
    c.view(torch.int64)[0] = 100
    return c


@pytest.mark.parametrize("opt_name", ["adam", "sgd"])
def test_static_entry_point_matches_the_caller_scratch_one_twice_in_a_row(opt_name):
    from distributedtensorflow_b200.ops import _lib
    lib = _lib.lib()
    L = _setup(lib)
    opt = {"adam": L.OPT_ADAM, "sgd": L.OPT_SGD}[opt_name]
    shapes = [(3136, 64), (1024, 10), (10,), (4099,)]
    ref_state = _state(shapes, seed=21)
    new_state = _clone(ref_state)
    ref_ctrl, new_ctrl = _ctrl(), _ctrl()
    ref_step = torch.zeros(1, dtype=torch.int64, device="cuda")
    new_step = torch.zeros(1, dtype=torch.int64, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")   # the scratch the old entry point needs
    hyper = L.ApplyHyper()
    hyper.lr, hyper.beta1, hyper.beta2, hyper.epsilon = 0.01, 0.9, 0.999, 1e-8
    hyper.momentum, hyper.grad_scale, hyper.use_nesterov, hyper.atomic_sgd = 0.0, 1.0, 0, 0
    stream = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device="cuda").manual_seed(5)
    for call in range(2):
        grads = [torch.randn(s["n"], device="cuda", generator=gen) for s in ref_state]
        torch.cuda.synchronize()
        rc = lib.dtf_push_apply(opt, _items(L, ref_state, grads), len(shapes), C.byref(hyper),
                                ref_ctrl.data_ptr(), ref_ctrl.data_ptr(), 1, 3, ref_step.data_ptr(),
                                counter.data_ptr(), stream)
        assert rc == 0, lib.dtf_last_error()
        rc = lib.dtf_push_apply_static(opt, _items(L, new_state, grads), len(shapes), C.byref(hyper),
                                       new_ctrl.data_ptr(), new_ctrl.data_ptr(), 1, 3, new_step.data_ptr(),
                                       stream)
        assert rc == 0, lib.dtf_last_error()
        torch.cuda.synchronize()
        for a, b in zip(ref_state, new_state):
            for key in ("var", "m", "v"):
                assert (a[key] - b[key]).abs().max().item() <= 1e-6, (call, key)
            assert (a["shadow"].float() - b["shadow"].float()).abs().max().item() <= 1e-2, call
        assert new_step.item() == ref_step.item() == 101 + call          # _finish ran exactly once
        assert torch.equal(ref_ctrl, new_ctrl)                           # step, beta powers, heartbeat
    assert (ref_state[0]["var"] - _state(shapes, seed=21)[0]["var"]).abs().max().item() > 0  # it did update


def test_static_entry_point_rejects_oversized_requests():
    from distributedtensorflow_b200.ops import _lib
    lib = _lib.lib()
    L = _setup(lib)
    state = _state([(16,)], seed=1)
    grads = [torch.zeros(16, device="cuda")]
    ctrl = _ctrl()
    hyper = L.ApplyHyper()
    hyper.lr, hyper.grad_scale = 0.1, 1.0
    items = _items(L, state, grads)
    stream = torch.cuda.current_stream().cuda_stream
    too_many = (L.ApplyItem * (L.MAX_APPLY_ITEMS + 1))()
    for k in range(L.MAX_APPLY_ITEMS + 1):
        too_many[k] = items[0]
    rc = lib.dtf_push_apply_static(L.OPT_SGD, too_many, L.MAX_APPLY_ITEMS + 1, C.byref(hyper), None,
                                   ctrl.data_ptr(), 1, 0, None, stream)
    assert rc != 0 and b"limit" in lib.dtf_last_error()
    items[0].n = (1 << 28) + 1                                     # beyond the documented maximum
    before = state[0]["var"].clone()
    rc = lib.dtf_push_apply_static(L.OPT_SGD, items, 1, C.byref(hyper), None, ctrl.data_ptr(), 1, 0, None, stream)
    assert rc != 0 and b"elements" in lib.dtf_last_error()
    torch.cuda.synchronize()
    assert torch.equal(before, state[0]["var"]) and ctrl.view(torch.int64)[0].item() == 100  # nothing ran
