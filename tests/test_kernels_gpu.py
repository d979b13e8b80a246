"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of the
same op. All tests need a B200 (``-m gpu``)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from distributedtensorflow_b200.ops import kernels
    from distributedtensorflow_b200.ops import _lib
    _lib.lib()  # fail loudly if the extension is missing on a GPU box
    return kernels


def _bf16(t):
    return t.to(torch.bfloat16)


def _rel_err(got, ref):
    got, ref = got.float(), ref.float()
    return (got - ref).abs().max().item() / max(ref.abs().max().item(), 1e-6)


def _padded(rows, cols, dtype=torch.bfloat16, fill=None):
    from distributedtensorflow_b200.ops.kernels import pad8
    t = torch.zeros(rows, pad8(cols), dtype=dtype, device="cuda")
    if fill is not None:
        t[:, :cols] = fill.to(dtype)
    return t


# ------------------------------------------------------------------------------------------------
# GEMM: all four operand layouts, ragged shapes, fused epilogues
# ------------------------------------------------------------------------------------------------
GEMM_SHAPES = [
    (128, 128, 64), (128, 64, 128), (256, 256, 512), (100, 100, 784), (100, 10, 100),
    (128, 1024, 3136), (128, 10, 1024), (784, 100, 100), (3136, 1024, 128), (384, 192, 200),
]


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K_", GEMM_SHAPES)
def test_gemm_layouts(K, M, N, K_, a_mn, b_mn):
    torch.manual_seed(M * 7 + N * 3 + K_)
    A = torch.randn(M, K_, device="cuda")
    Bm = torch.randn(N, K_, device="cuda")
    a_store = _padded(K_, M, fill=A.t()) if a_mn else _padded(M, K_, fill=A)
    b_store = _padded(K_, N, fill=Bm.t()) if b_mn else _padded(N, K_, fill=Bm)
    out = torch.full((M, N), float("nan"), device="cuda")
    K.gemm(a_store, b_store, out, M, N, K_, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    ref = _bf16(A).float() @ _bf16(Bm).float().t()
    err = _rel_err(out, ref)
    assert err < 2e-3, f"rel err {err}"


@pytest.mark.parametrize("block_n", [64, 128])
def test_gemm_bias_relu_bf16_out(K, block_n):
    torch.manual_seed(0)
    M, N, K_ = 128, 200, 320
    A = torch.randn(M, K_, device="cuda")
    W = torch.randn(K_, N, device="cuda") * 0.1
    bias = torch.randn(N, device="cuda")
    a = _padded(M, K_, fill=A)
    w = _padded(K_, N, fill=W)  # [K, N] row-major == MN-major B
    out = torch.full((M, w.shape[1]), 7.0, dtype=torch.bfloat16, device="cuda")
    K.gemm(a, w, out, M, N, K_, b_mn=True, bias=bias, relu=True, block_n=block_n)
    torch.cuda.synchronize()
    ref = torch.relu(_bf16(A).float() @ _bf16(W).float() + bias)
    assert _rel_err(out[:, :N], ref) < 1e-2
    assert (out[:, N:] == 0).all(), "row padding must be written as zeros"


# persistent CTA-pair kernel (tcgen05.mma.cta_group::2, 256 x 256 tiles, two TMEM accumulators)
PAIR_SHAPES = [(512, 512, 256), (256, 256, 64), (300, 520, 200), (1024, 2048, 1000),
               (4096, 6144, 128)]   # last: 384 tiles on 74 pairs -> both accumulator phases wrap


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K_", PAIR_SHAPES)
def test_gemm_pair_layouts(K, M, N, K_, a_mn, b_mn):
    torch.manual_seed(M + N * 3 + K_)
    A = torch.randn(M, K_, device="cuda")
    Bm = torch.randn(N, K_, device="cuda")
    a_store = _padded(K_, M, fill=A.t()) if a_mn else _padded(M, K_, fill=A)
    b_store = _padded(K_, N, fill=Bm.t()) if b_mn else _padded(N, K_, fill=Bm)
    out = torch.full((M, N), float("nan"), device="cuda")
    K.gemm(a_store, b_store, out, M, N, K_, a_mn=a_mn, b_mn=b_mn, block_n=256)
    torch.cuda.synchronize()
    ref = _bf16(A).float() @ _bf16(Bm).float().t()
    err = _rel_err(out, ref)
    assert err < 2e-3, f"rel err {err}"
    # element-wise too: a swapped B half or accumulator stage would keep the norm but move values
    assert torch.allclose(out, ref, rtol=2e-2, atol=2e-2 * ref.abs().max().item())


def test_gemm_pair_fused_epilogues(K):
    torch.manual_seed(5)
    M, N, K_ = 520, 776, 328
    A = torch.randn(M, K_, device="cuda")
    W = torch.randn(K_, N, device="cuda") * 0.1
    bias = torch.randn(N, device="cuda")
    a, w = _padded(M, K_, fill=A), _padded(K_, N, fill=W)
    out = torch.full((M, w.shape[1]), 7.0, dtype=torch.bfloat16, device="cuda")
    K.gemm(a, w, out, M, N, K_, b_mn=True, bias=bias, relu=True, block_n=256)
    torch.cuda.synchronize()
    ref = torch.relu(_bf16(A).float() @ _bf16(W).float() + bias)
    assert _rel_err(out[:, :N], ref) < 1e-2
    # ReLU-gradient mask + fp32 accumulate-into (red.add) epilogues
    dY = torch.randn(M, K_, device="cuda")
    Wt = torch.randn(N, K_, device="cuda")
    act = torch.randn(M, N, device="cuda")
    mask = _padded(M, N, fill=torch.relu(act))
    out2 = torch.zeros(M, mask.shape[1], dtype=torch.bfloat16, device="cuda")
    K.gemm(_padded(M, K_, fill=dY), _padded(N, K_, fill=Wt), out2, M, N, K_, mask=mask, block_n=256)
    acc = torch.ones(M, N, device="cuda")
    K.gemm(_padded(M, K_, fill=dY), _padded(N, K_, fill=Wt), acc, M, N, K_, reduce=True, alpha=0.5, block_n=256)
    torch.cuda.synchronize()
    ref2 = _bf16(dY).float() @ _bf16(Wt).float().t()
    assert _rel_err(out2[:, :N], ref2 * (act > 0)) < 1e-2
    assert _rel_err(acc, 1.0 + 0.5 * ref2) < 2e-3


def test_gemm_relu_grad_mask(K):
    torch.manual_seed(1)
    M, N, K_ = 128, 104, 16
    dY = torch.randn(M, K_, device="cuda")
    W = torch.randn(N, K_, device="cuda")
    act = torch.randn(M, N, device="cuda")
    a = _padded(M, K_, fill=dY)
    w = _padded(N, K_, fill=W)
    mask = _padded(M, N, fill=torch.relu(act))
    out = torch.zeros(M, mask.shape[1], dtype=torch.bfloat16, device="cuda")
    K.gemm(a, w, out, M, N, K_, mask=mask)
    torch.cuda.synchronize()
    ref = (_bf16(dY).float() @ _bf16(W).float().t()) * (act > 0)
    assert _rel_err(out[:, :N], ref) < 1e-2


def test_gemm_split_k_reduce(K):
    torch.manual_seed(2)
    M, N, K_ = 256, 64, 4096
    A = torch.randn(M, K_, device="cuda")
    Bm = torch.randn(N, K_, device="cuda")
    a, b = _padded(M, K_, fill=A), _padded(N, K_, fill=Bm)
    out = torch.zeros(M, N, device="cuda")
    K.gemm(a, b, out, M, N, K_, reduce=True, splits=8, alpha=0.5)
    torch.cuda.synchronize()
    ref = 0.5 * (_bf16(A).float() @ _bf16(Bm).float().t())
    assert _rel_err(out, ref) < 2e-3


@pytest.mark.parametrize("M,N,K_,splits,block_n", [(128, 1024, 3136, 9, 64), (128, 10, 1024, 8, 64),
                                                   (128, 3136, 1024, 3, 64), (100, 200, 784, 4, 128)])
def test_gemm_split_k_fixup_keeps_fused_epilogue(K, M, N, K_, splits, block_n):
    """Split-K with the in-kernel fix-up: K-slices sum into an fp32 workspace, the last slice of
    each tile applies bias + ReLU and writes bf16; the tile counters re-arm themselves."""
    torch.manual_seed(12)
    A = torch.randn(M, K_, device="cuda")
    W = torch.randn(K_, N, device="cuda") * 0.05
    bias = torch.randn(N, device="cuda")
    a, w = _padded(M, K_, fill=A), _padded(K_, N, fill=W)
    ldws = (N + 3) // 4 * 4
    tiles = ((M + 127) // 128) * ((N + block_n - 1) // block_n)
    counters = torch.zeros(tiles, dtype=torch.int32, device="cuda")
    ref = torch.relu(_bf16(A).float() @ _bf16(W).float() + bias)
    for _ in range(2):                                   # second run: counters must be back at zero
        ws = torch.zeros(M, ldws, device="cuda")
        out = torch.full((M, w.shape[1]), 7.0, dtype=torch.bfloat16, device="cuda")
        K.gemm(a, w, out, M, N, K_, b_mn=True, bias=bias, relu=True, splits=splits, block_n=block_n,
               ws=ws, counters=counters)
        torch.cuda.synchronize()
        assert _rel_err(out[:, :N], ref) < 1e-2
        assert (out[:, N:] == 0).all()
        assert (counters == 0).all()


@pytest.mark.parametrize("opt", ["adam", "sgd", "momentum", "adagrad"])
@pytest.mark.parametrize("M,N,B,slices", [(3136, 1024, 128, 1), (1024, 10, 128, 1), (784, 104, 100, 3)])
def test_gemm_fused_push_apply_matches_gemm_then_apply(K, opt, M, N, B, slices):
    """dW = X^T dY with the optimizer applied in the epilogue (no gradient buffer) == the same GEMM
    writing fp32 gradients followed by dtf_push_apply; the variable may be cut into row slices."""
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(14)
    code = {"sgd": L.OPT_SGD, "momentum": L.OPT_MOMENTUM, "adagrad": L.OPT_ADAGRAD, "adam": L.OPT_ADAM}[opt]
    X = _padded(B, M, fill=torch.randn(B, M, device="cuda"))          # [K=B, M]  (MN-major A)
    dY = _padded(B, N, fill=torch.randn(B, N, device="cuda") * 0.1)   # [K=B, N]  (MN-major B)
    ld = K.pad8(N)
    var0 = torch.randn(M * N, device="cuda")
    m0 = torch.rand(M * N, device="cuda") * 0.1 + 0.1
    v0 = torch.rand(M * N, device="cuda") * 0.01
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")
    ctrl.view(torch.float32)[2], ctrl.view(torch.float32)[3] = 0.9 ** 2, 0.999 ** 2
    hyper = K.make_hyper(0.01, momentum=0.9)
    # reference: GEMM -> gradient buffer -> push_apply
    grad = torch.zeros(M, N, device="cuda")
    K.gemm(X, dY, grad, M, N, B, a_mn=True, b_mn=True)
    rv, rm, rvv = var0.clone(), m0.clone(), v0.clone()
    rsh = torch.zeros(M, ld, dtype=torch.bfloat16, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    items = K.make_apply_items([dict(var=rv.data_ptr(), slot0=rm.data_ptr(), slot1=rvv.data_ptr(), shadow=rsh.data_ptr(),
                                     grad=grad.data_ptr(), n=M * N, cols=N, shadow_ld=ld)])
    K.push_apply(code, items, 1, hyper, ctrl.data_ptr(), ctrl.data_ptr(), False, 0, None, counter.data_ptr())
    # fused: separate allocations per row slice, as on a flat-sliced cluster
    rows = -(-M // slices)
    rows = -(-rows // 64) * 64 if slices > 1 else M
    sl = []
    keep = []
    for i in range(slices):
        b, e = min(M, i * rows), min(M, (i + 1) * rows)
        fv, fm, fvv = var0[b * N:e * N].clone(), m0[b * N:e * N].clone(), v0[b * N:e * N].clone()
        fsh = torch.zeros(max(e - b, 1), ld, dtype=torch.bfloat16, device="cuda")
        keep.append((b, e, fv, fm, fvv, fsh))
        sl.append(dict(var=fv.data_ptr(), slot0=fm.data_ptr() if opt != "sgd" else None,
                       slot1=fvv.data_ptr() if opt == "adam" else None, shadow=fsh.data_ptr()))
    K.gemm(X, dY, None, M, N, B, a_mn=True, b_mn=True,
           apply=dict(opt=code, hyper=hyper, rows_per_slice=rows, shadow_ld=ld, ctrl_powers=ctrl.data_ptr(), slices=sl))
    torch.cuda.synchronize()
    for b, e, fv, fm, fvv, fsh in keep:
        assert (fv - rv[b * N:e * N]).abs().max().item() < 1e-6
        if opt != "sgd":
            assert (fm - rm[b * N:e * N]).abs().max().item() < 1e-6
        if opt == "adam":
            assert (fvv - rvv[b * N:e * N]).abs().max().item() < 1e-6
        # serving copies agree to one bf16 ulp (the two kernels contract FMAs differently)
        assert (fsh[:e - b, :N].float() - rsh[b:e, :N].float()).abs().max().item() <= 0.01 * rv.abs().max().item()
    assert (rv - var0).abs().max().item() > 0


def test_pdl_launches_give_identical_results(K):
    """Programmatic dependent launch only overlaps prologues: results must not change."""
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(13)
    M, N, K_ = 128, 256, 512
    a = _padded(M, K_, fill=torch.randn(M, K_, device="cuda"))
    b = _padded(N, K_, fill=torch.randn(N, K_, device="cuda"))
    outs = []
    for pdl in (0, 1):
        old = L.lib().dtf_set_pdl(pdl)
        o1 = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        o2 = torch.zeros(M, M, device="cuda")
        cs = torch.zeros(N, device="cuda")
        for _ in range(3):                               # a chain of dependent kernels
            K.gemm(a, b, o1, M, N, K_)
            K.gemm(o1, o1, o2, M, M, N)
            K.fill_u32(cs, 0)
            K.colsum_bf16(o1, M, N, cs)
        torch.cuda.synchronize()
        L.lib().dtf_set_pdl(old)
        outs.append((o1.clone(), o2.clone(), cs.clone()))
    (a1, a2, acs), (b1, b2, bcs) = outs
    assert torch.equal(a1, b1) and torch.equal(a2, b2)
    assert (acs - bcs).abs().max().item() < 1e-3        # column sums use float atomics


# ------------------------------------------------------------------------------------------------
# loss
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,Cn", [(128, 10), (100, 10), (37, 33)])
def test_softmax_xent_sparse(K, B, Cn):
    torch.manual_seed(3)
    logits = torch.randn(B, Cn, device="cuda") * 3
    labels = torch.randint(0, Cn, (B,), device="cuda", dtype=torch.int32)
    loss = torch.zeros(1, device="cuda")
    dl = _padded(B, Cn)
    db = torch.zeros(Cn, device="cuda")
    K.softmax_xent_sparse(logits, labels, loss, dl, db)
    torch.cuda.synchronize()
    lref = logits.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lref, labels.long())
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-4 * max(1.0, abs(ref.item()))
    assert (dl[:, :Cn].float() - lref.grad).abs().max().item() < 1e-2 / B + 1e-4
    assert (db - lref.grad.sum(0)).abs().max().item() < 2e-2
    assert (dl[:, Cn:] == 0).all()


def test_softmax_xent_dense_clipped_sum(K):
    torch.manual_seed(4)
    B, Cn = 100, 10
    logits = torch.randn(B, Cn, device="cuda") * 4
    logits[0, 3] = 60.0  # forces probabilities below the 1e-10 clip for the other classes
    labels = torch.nn.functional.one_hot(torch.randint(0, Cn, (B,), device="cuda"), Cn).float()
    loss = torch.zeros(1, device="cuda")
    dl = _padded(B, Cn)
    K.softmax_xent_dense(logits, labels, loss, dl)
    torch.cuda.synchronize()
    lref = logits.clone().requires_grad_(True)
    y = torch.softmax(lref, dim=1)
    ref = -(labels * torch.log(torch.clamp(y, 1e-10, 1.0))).sum()
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-3 * max(1.0, abs(ref.item()))
    assert (dl[:, :Cn].float() - lref.grad).abs().max().item() < 2e-2


# ------------------------------------------------------------------------------------------------
# push + apply (TF1 update rules, fp64 restatement as oracle)
# ------------------------------------------------------------------------------------------------
def _apply_ref(opt, var, s0, s1, g, h, b1p, b2p):
    var, s0, s1, g = (t.double() for t in (var, s0, s1, g))
    if opt == "sgd":
        var = var - h["lr"] * g
    elif opt == "momentum":
        s0 = s0 * h["momentum"] + g
        var = var - h["lr"] * s0
    elif opt == "adagrad":
        s0 = s0 + g * g
        var = var - h["lr"] * g / s0.sqrt()
    else:
        lr_t = h["lr"] * math.sqrt(1 - b2p) / (1 - b1p)
        s0 = s0 + (g - s0) * (1 - h["beta1"])
        s1 = s1 + (g * g - s1) * (1 - h["beta2"])
        var = var - lr_t * s0 / (s1.sqrt() + h["epsilon"])
    return var, s0, s1


@pytest.mark.parametrize("opt", ["sgd", "momentum", "adagrad", "adam"])
def test_push_apply_matches_tf1_rules(K, opt):
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(5)
    code = {"sgd": L.OPT_SGD, "momentum": L.OPT_MOMENTUM, "adagrad": L.OPT_ADAGRAD, "adam": L.OPT_ADAM}[opt]
    shapes = [(3136, 1024), (1024,), (1024, 10), (10,), (5, 5, 1, 32), (4099,)]
    h = dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, momentum=0.9)
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")  # 256-byte control block
    ctrl_f = ctrl.view(torch.float32)
    b1p, b2p = 0.9 ** 3, 0.999 ** 3
    ctrl_f[2], ctrl_f[3] = b1p, b2p
    ctrl.view(torch.int64)[0] = 41
    step_out = torch.zeros(1, dtype=torch.int64, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    entries, refs, keep = [], [], []
    for shp in shapes:
        n = math.prod(shp)
        cols = shp[-1]
        ld = K.pad8(cols)
        var = torch.randn(n, device="cuda")
        s0 = torch.rand(n, device="cuda") * (0.1 if opt != "adagrad" else 1.0) + (0.1 if opt == "adagrad" else 0)
        s1 = torch.rand(n, device="cuda") * 0.01
        g = torch.randn(n, device="cuda")
        shadow = torch.zeros(n // cols, ld, dtype=torch.bfloat16, device="cuda")
        refs.append(_apply_ref(opt, var, s0, s1, g, h, b1p, b2p))
        keep.append((var, s0, s1, g, shadow, cols))
        entries.append(dict(var=var.data_ptr(), slot0=s0.data_ptr(), slot1=s1.data_ptr(),
                            shadow=shadow.data_ptr(), grad=g.data_ptr(), n=n, first=0, cols=cols,
                            shadow_ld=ld))
    items = K.make_apply_items(entries)
    hyper = K.make_hyper(h["lr"], h["beta1"], h["beta2"], h["epsilon"], h["momentum"])
    K.push_apply(code, items, len(entries), hyper, ctrl.data_ptr(), ctrl.data_ptr(), True, 2,
                 step_out.data_ptr(), counter.data_ptr())
    torch.cuda.synchronize()
    for (var, s0, s1, g, shadow, cols), (rv, r0, r1) in zip(keep, refs):
        assert (var.double() - rv).abs().max().item() < 1e-5
        if opt != "sgd":
            assert (s0.double() - r0).abs().max().item() < 1e-5
        if opt == "adam":
            assert (s1.double() - r1).abs().max().item() < 1e-5
        assert (shadow[:, :cols].reshape(-1).float() - var).abs().max().item() < 2e-2
    assert ctrl.view(torch.int64)[0].item() == 42 and step_out.item() == 42
    assert counter.item() == 0, "last-CTA election counter must wrap back to zero"
    if opt == "adam":
        assert abs(ctrl_f[2].item() - b1p * 0.9) < 1e-7 and abs(ctrl_f[3].item() - b2p * 0.999) < 1e-7
    hb = ctrl.view(torch.int64)[4 + 2].item()
    assert hb == 42


def test_push_apply_atomic_sgd_is_exact_sum(K):
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(6)
    n = 100003
    var = torch.zeros(n, device="cuda")
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    grads = [torch.randn(n, device="cuda") for _ in range(4)]
    hyper = K.make_hyper(0.5, atomic_sgd=True)
    streams = [torch.cuda.Stream() for _ in grads]
    torch.cuda.synchronize()
    for g, s in zip(grads, streams):  # concurrent writers on different streams
        with torch.cuda.stream(s):
            items = K.make_apply_items([dict(var=var.data_ptr(), grad=g.data_ptr(), n=n, cols=1)])
            K.push_apply(L.OPT_SGD, items, 1, hyper, None, ctrl.data_ptr(), False, 0, None,
                         counter.data_ptr())
    torch.cuda.synchronize()
    ref = -0.5 * torch.stack(grads).double().sum(0)
    assert (var.double() - ref).abs().max().item() < 1e-5


def test_pull_bulk_and_refresh_shadow(K):
    torch.manual_seed(7)
    srcs = [torch.randn(n, device="cuda") for n in (4, 1024, 3136 * 1024, 12)]
    dsts = [torch.zeros_like(s) for s in srcs]
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")
    ctrl.view(torch.int64)[0] = 1234
    pulled = torch.zeros(1, dtype=torch.int64, device="cuda")
    K.pull_bulk([(s.data_ptr(), d.data_ptr(), s.numel() * 4) for s, d in zip(srcs, dsts)],
                ctrl.data_ptr(), pulled.data_ptr())
    torch.cuda.synchronize()
    for s, d in zip(srcs, dsts):
        assert torch.equal(s, d)
    assert pulled.item() == 1234
    var = torch.randn(1024 * 10, device="cuda")
    shadow = torch.zeros(1024, 16, dtype=torch.bfloat16, device="cuda")
    K.refresh_shadow(var.data_ptr(), shadow.data_ptr(), var.numel(), 0, 10, 16)
    torch.cuda.synchronize()
    assert torch.equal(shadow[:, :10], var.view(1024, 10).to(torch.bfloat16))


def test_elementwise_helpers(K):
    torch.manual_seed(8)
    x = torch.randint(0, 256, (100, 784), dtype=torch.uint8, device="cuda")
    dst = _padded(100, 784)
    K.u8_to_bf16(x, dst)
    f = torch.randn(33, 10, device="cuda")
    fd = _padded(33, 10)
    K.f32_to_bf16(f, fd)
    m = _padded(300, 70, fill=torch.randn(300, 70, device="cuda"))
    cs = torch.zeros(70, device="cuda")
    K.colsum_bf16(m, 300, 70, cs)
    z = torch.ones(1000, device="cuda")
    K.fill_u32(z, 0)
    torch.cuda.synchronize()
    assert torch.equal(dst[:, :784], (x.float() * (1.0 / 255.0)).to(torch.bfloat16))
    assert torch.equal(fd[:, :10], f.to(torch.bfloat16)) and (fd[:, 10:] == 0).all()
    assert (cs - m[:, :70].float().sum(0)).abs().max().item() < 1e-3
    assert (z == 0).all()


# ------------------------------------------------------------------------------------------------
# CNN layers (run_mnist_distributed.py:52-65)
# ------------------------------------------------------------------------------------------------
def _conv1_ref(x_u8, w1, b1):
    x = (x_u8.float() / 255.0).view(-1, 1, 28, 28)
    w = w1.view(5, 5, 1, 32).permute(3, 2, 0, 1).contiguous()
    conv = torch.nn.functional.conv2d(x, w, b1, padding=2)
    return x, w, conv


@pytest.mark.parametrize("B", [128, 6])
def test_conv1_fwd_bwd(K, B):
    torch.manual_seed(9)
    x_u8 = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, device="cuda")
    w1 = torch.randn(25, 32, device="cuda") * 0.2
    b1 = torch.randn(32, device="cuda") * 0.1
    pool1p = torch.zeros(B, 18, 18, 32, dtype=torch.bfloat16, device="cuda")
    amax1 = torch.zeros(B, 14, 14, 32, dtype=torch.uint8, device="cuda")
    K.conv1_fwd(x_u8, w1, b1, pool1p, amax1, 1.0 / 255.0)
    torch.cuda.synchronize()
    x, w, conv = _conv1_ref(x_u8, w1, b1)
    w.requires_grad_(True)
    b = b1.clone().requires_grad_(True)
    conv = torch.nn.functional.conv2d(x, w, b, padding=2)
    pooled = torch.nn.functional.max_pool2d(torch.relu(conv), 2)
    got = pool1p[:, 2:16, 2:16, :].permute(0, 3, 1, 2).float()
    assert (got - pooled).abs().max().item() < 2e-2 * max(1.0, pooled.abs().max().item())
    assert (pool1p[:, :2] == 0).all() and (pool1p[:, :, 16:] == 0).all()
    # backward: random upstream gradient
    dpool1 = (torch.randn(B, 14, 14, 32, device="cuda") * 0.1).to(torch.bfloat16)
    dw = torch.zeros(25, 32, device="cuda")
    db = torch.zeros(32, device="cuda")
    K.conv1_bwd(x_u8, dpool1, pool1p, amax1, dw, db, 1.0 / 255.0)
    torch.cuda.synchronize()
    pooled.backward(dpool1.float().permute(0, 3, 1, 2))
    dw_ref = w.grad.permute(2, 3, 1, 0).reshape(25, 32)
    assert _rel_err(dw, dw_ref) < 2e-2
    assert _rel_err(db, b.grad) < 2e-2


def _conv2_setup(B):
    torch.manual_seed(10 + B)
    a = torch.relu(torch.randn(B, 14, 14, 32, device="cuda"))
    pool1p = torch.zeros(B, 18, 18, 32, dtype=torch.bfloat16, device="cuda")
    pool1p[:, 2:16, 2:16, :] = a.to(torch.bfloat16)
    w2 = (torch.randn(5, 5, 32, 64, device="cuda") * 0.05).to(torch.bfloat16)
    b2 = torch.randn(64, device="cuda") * 0.1
    return pool1p, w2, b2


@pytest.mark.parametrize("halo", [True, False])
@pytest.mark.parametrize("B", [128, 4, 6])
def test_conv2_fwd(K, B, halo):
    pool1p, w2, b2 = _conv2_setup(B)
    pool2 = torch.zeros(B, 7, 7, 64, dtype=torch.bfloat16, device="cuda")
    amax2 = torch.zeros(B, 7, 7, 64, dtype=torch.uint8, device="cuda")
    K.conv2_fwd(pool1p, w2.data_ptr(), b2, pool2, amax2, halo=halo)
    torch.cuda.synchronize()
    x = pool1p[:, 2:16, 2:16, :].float().permute(0, 3, 1, 2)
    w = w2.float().permute(3, 2, 0, 1)
    conv = torch.nn.functional.conv2d(x, w, b2, padding=2)
    pooled, idx = torch.nn.functional.max_pool2d(torch.relu(conv), 2, return_indices=True)
    got = pool2.float().permute(0, 3, 1, 2)
    err = (got - pooled).abs().max().item()
    assert err < 2e-2 * max(1.0, pooled.abs().max().item()), f"abs err {err}"
    # argmax agrees wherever the window has a unique, clearly separated maximum
    convw = conv.view(B, 64, 7, 2, 7, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, 64, 7, 7, 4)
    top2 = convw.topk(2, dim=-1).values
    clear = (top2[..., 0] - top2[..., 1]) > 0.05
    ref_am = convw.argmax(-1)
    got_am = amax2.permute(0, 3, 1, 2).long()
    assert (got_am[clear] == ref_am[clear]).all()


@pytest.mark.parametrize("halo", [True, False])
@pytest.mark.parametrize("B", [128, 6])
def test_conv2_backward(K, B, halo):
    pool1p, w2, b2 = _conv2_setup(B)
    # upstream gradient w.r.t. the conv2 pre-activation, in the zero-bordered layout
    dy = (torch.randn(B, 14, 14, 64, device="cuda") * 0.1).to(torch.bfloat16)
    dconv2p = torch.zeros(B, 18, 18, 64, dtype=torch.bfloat16, device="cuda")
    dconv2p[:, 2:16, 2:16, :] = dy
    x = pool1p[:, 2:16, 2:16, :].float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    w = w2.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    conv = torch.nn.functional.conv2d(x, w, None, padding=2)
    conv.backward(dy.float().permute(0, 3, 1, 2))
    dw = torch.zeros(800, 64, device="cuda")
    K.conv2_wgrad(pool1p, dconv2p, dw)
    dpool1 = torch.zeros(B, 14, 14, 32, dtype=torch.bfloat16, device="cuda")
    K.conv2_dgrad(dconv2p, w2.data_ptr(), dpool1, halo=halo)
    torch.cuda.synchronize()
    dw_ref = w.grad.permute(2, 3, 1, 0).reshape(800, 64)
    assert _rel_err(dw, dw_ref) < 1e-2, f"wgrad rel err {_rel_err(dw, dw_ref)}"
    dx_ref = x.grad.permute(0, 2, 3, 1)
    assert _rel_err(dpool1, dx_ref) < 2e-2, f"dgrad rel err {_rel_err(dpool1, dx_ref)}"


def test_pool2_bwd(K):
    torch.manual_seed(11)
    B = 16
    conv = torch.randn(B, 64, 14, 14, device="cuda", requires_grad=True)
    pooled, idx = torch.nn.functional.max_pool2d(torch.relu(conv), 2, return_indices=True)
    # our argmax encoding: position inside the 2x2 window
    convw = conv.detach().view(B, 64, 7, 2, 7, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, 64, 7, 7, 4)
    amax2 = convw.argmax(-1).permute(0, 2, 3, 1).contiguous().to(torch.uint8)
    pool2 = pooled.detach().permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
    d = (torch.randn(B, 7, 7, 64, device="cuda")).to(torch.bfloat16)
    dpool2 = d.reshape(B, 3136).contiguous()
    dconv2p = torch.zeros(B, 18, 18, 64, dtype=torch.bfloat16, device="cuda")
    db2 = torch.zeros(64, device="cuda")
    K.pool2_bwd(dpool2, pool2, amax2, dconv2p, db2)
    torch.cuda.synchronize()
    pooled.backward(d.float().permute(0, 3, 1, 2))
    ref = conv.grad.permute(0, 2, 3, 1)
    got = dconv2p[:, 2:16, 2:16, :].float()
    assert (got - ref).abs().max().item() < 1e-2
    assert (db2 - ref.sum((0, 1, 2))).abs().max().item() < 0.05


def test_optimizer_rules_division_and_sqrt_track_ieee_rounding():
    """The optimizer kernels are built without fast-math but use a branch-free division / sqrt
    (apply_rules.cuh): on operands in the range optimizer state lives in they must agree with IEEE
    round-to-nearest - bit-identical almost everywhere, never more than one ulp away."""
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(0)
    n = 1 << 22
    dev = "cuda"
    # numerators like lr_t * m, denominators like sqrt(v) + eps: 10^-12 .. 10^3, both signs upstairs
    a = (torch.randn(n, device=dev) * torch.pow(10.0, torch.empty(n, device=dev).uniform_(-12, 3))).float()
    b = torch.pow(10.0, torch.empty(n, device=dev).uniform_(-12, 3)).float()
    b[:4] = torch.tensor([0.0, 1e-8, 1.0, 4.0], device=dev)
    out = [torch.empty(n, device=dev) for _ in range(4)]
    rc = L.lib().dtf_debug_div_sqrt(a.data_ptr(), b.data_ptr(), *[o.data_ptr() for o in out], n, L.current_stream())
    L.check(rc, "dtf_debug_div_sqrt")
    torch.cuda.synchronize()
    qf, qi, sf, si = out
    assert sf[0].item() == 0.0 and sf[2].item() == 1.0 and sf[3].item() == 2.0
    for fast, ieee, name in ((qf[1:], qi[1:], "div"), (sf, si, "sqrt")):
        ulp = (fast.view(torch.int32) - ieee.view(torch.int32)).abs()
        assert ulp.max().item() <= 1, (name, ulp.max().item())
        assert (ulp != 0).float().mean().item() < 1e-3, (name, (ulp != 0).float().mean().item())


@pytest.mark.parametrize("B,splits", [(128, 1), (128, 3), (37, 3), (200, 2)])
def test_gemm_pool_scatter_epilogue_equals_gemm_plus_pool2_bwd(K, B, splits):
    """The dX GEMM with conv2's ReLU + max-pool gradient as its epilogue against the two-kernel form
    (GEMM -> dpool2 -> pool2_bwd) and against PyTorch autograd on the same operands."""
    torch.manual_seed(B + splits)
    dev = "cuda"
    Kd = 256
    dz = (torch.randn(B, Kd, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(3136, Kd, device=dev) * 0.1).to(torch.bfloat16)          # [N, K], K contiguous
    conv = torch.randn(B, 64, 14, 14, device=dev, requires_grad=True)
    pooled, _ = torch.nn.functional.max_pool2d(torch.relu(conv), 2, return_indices=True)
    convw = conv.detach().view(B, 64, 7, 2, 7, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, 64, 7, 7, 4)
    amax2 = convw.argmax(-1).permute(0, 2, 3, 1).contiguous().to(torch.uint8)
    pool2 = pooled.detach().permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).reshape(B, 3136)
    ws = torch.zeros(B, 3136, device=dev)
    ct = torch.zeros(256, dtype=torch.int32, device=dev)          # one counter per 128 x 64 output tile
    kw = dict(block_n=64)
    if splits > 1:
        kw.update(splits=splits, ws=ws, counters=ct)
    # two kernels
    dpool2 = torch.full((B, 3136), float("nan"), dtype=torch.bfloat16, device=dev)
    K.gemm(dz, w, dpool2, B, 3136, Kd, **kw)
    ref_out = torch.zeros(B, 18, 18, 64, dtype=torch.bfloat16, device=dev)
    ref_db = torch.zeros(64, device=dev)
    K.pool2_bwd(dpool2, pool2.view(B, 7, 7, 64), amax2, ref_out, ref_db)
    # one kernel
    ws.zero_()
    got_out = torch.zeros(B, 18, 18, 64, dtype=torch.bfloat16, device=dev)
    got_out[:, 2:16, 2:16, :] = float("nan")                       # every interior element must be written
    got_db = torch.zeros(64, device=dev)
    K.gemm(dz, w, None, B, 3136, Kd, mask=pool2, pool=dict(out=got_out, amax=amax2, dbias=got_db), **kw)
    torch.cuda.synchronize()
    assert torch.isfinite(got_out.float()).all()
    assert got_out[:, :2].abs().sum() == 0 and got_out[:, 16:].abs().sum() == 0       # the zero border is untouched
    assert got_out[:, :, :2].abs().sum() == 0 and got_out[:, :, 16:].abs().sum() == 0
    d = (got_out.float() - ref_out.float()).abs()
    assert d.max().item() <= 0.03 * ref_out.float().abs().max().item() + 1e-3        # one bf16 rounding vs two
    # autograd oracle
    g = (dz.float() @ w.float().t()).view(B, 7, 7, 64).permute(0, 3, 1, 2)
    pooled.backward(g)
    oracle = conv.grad.permute(0, 2, 3, 1)
    assert _rel_err(got_out[:, 2:16, 2:16, :], oracle) < 1e-2
    assert _rel_err(got_db, oracle.sum((0, 1, 2))) < 1e-2
    assert _rel_err(ref_db, oracle.sum((0, 1, 2))) < 2e-2


# ---------------------------------------------------------------------------------------------
# "owner applies" push: mailbox_push (worker) + mailbox_drain (owner), csrc/kernels/mailbox.cu
# ---------------------------------------------------------------------------------------------
def _adam_ref(var, m, v, g, lr, b1, b2, eps, b1p, b2p):
    lr_t = lr * math.sqrt(1.0 - b2p) / (1.0 - b1p)
    m = m + (g - m) * (1.0 - b1)
    v = v + (g * g - v) * (1.0 - b2)
    var = var - lr_t * m / (v.sqrt() + eps)
    return var, m, v


def test_mailbox_push_and_drain_apply_pending_gradients_in_worker_order(K):
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(11)
    dev = "cuda"
    W, depth = 3, 2
    rows, cols = 300, 40
    n0, n1 = rows * cols, 10                       # a matrix with a serving copy + a ragged bias
    off1 = (n0 * 4 + 255) // 256 * 256
    stride = off1 + 256                            # bytes per (worker, buffer) slot
    slots = torch.zeros(W * depth * stride // 4, device=dev)
    flags = torch.zeros(128, dtype=torch.int32, device=dev)
    todo = torch.zeros(32, dtype=torch.int32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    ctrl = torch.zeros(64, dtype=torch.int32, device=dev)
    lr, b1, b2, eps = 1e-2, 0.9, 0.999, 1e-8
    ctrl.view(torch.float32)[2], ctrl.view(torch.float32)[3] = b1, b2
    step_out = torch.zeros(1, dtype=torch.int64, device=dev)
    var = [torch.randn(n0, device=dev), torch.randn(n1, device=dev)]
    m = [torch.zeros(n0, device=dev), torch.zeros(n1, device=dev)]
    v = [torch.zeros(n0, device=dev), torch.zeros(n1, device=dev)]
    shadow = torch.zeros(rows, cols, dtype=torch.bfloat16, device=dev)
    ref = [[x.clone() for x in var], [x.clone() for x in m], [x.clone() for x in v]]
    b1p, b2p = b1, b2
    items = K.make_apply_items([
        dict(var=var[0].data_ptr(), slot0=m[0].data_ptr(), slot1=v[0].data_ptr(), shadow=shadow.data_ptr(),
             grad=slots.data_ptr(), n=n0, cols=cols, shadow_ld=cols),
        dict(var=var[1].data_ptr(), slot0=m[1].data_ptr(), slot1=v[1].data_ptr(), shadow=None,
             grad=slots.data_ptr() + off1, n=n1, cols=n1, shadow_ld=n1)])
    hyper = K.make_hyper(lr, b1, b2, eps)
    seq = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(W)]
    pcount = torch.zeros(4, dtype=torch.int32, device=dev)
    minc = torch.zeros(1, dtype=torch.int32, device=dev)

    def push(w, grads):
        g1 = torch.zeros(16, device=dev)
        g1[:n1] = grads[1]
        base = slots.data_ptr() + w * depth * stride
        K.mailbox_push([(grads[0].data_ptr(), base, n0 * 4, stride)], depth, w, [flags.data_ptr()], False,
                       seq[w].data_ptr(), None, None)
        K.mailbox_push([(g1.data_ptr(), base + off1, 16 * 4, stride)], depth, w, [flags.data_ptr()], True,
                       seq[w].data_ptr(), pcount.data_ptr(), minc.data_ptr())
        torch.cuda.synchronize()

    def drain():
        K.mailbox_drain(L.OPT_ADAM, items, 2, stride, depth, W, hyper, ctrl.data_ptr(), ctrl.data_ptr(),
                        flags.data_ptr(), todo.data_ptr(), counter.data_ptr(), step_out.data_ptr())
        torch.cuda.synchronize()

    total = 0
    for rnd, pushers in enumerate([[0, 2], [0, 1, 2], [], [1]]):
        for w in pushers:
            g = [torch.randn(n0, device=dev), torch.randn(n1, device=dev)]
            push(w, g)
            for t in range(2):
                ref[0][t], ref[1][t], ref[2][t] = _adam_ref(ref[0][t], ref[1][t], ref[2][t], g[t],
                                                            lr, b1, b2, eps, b1p, b2p)
            b1p, b2p = b1p * b1, b2p * b2
        drain()
        total += len(pushers)
        assert int(step_out.item()) == total
        assert ctrl.view(torch.int64)[0].item() == total
        for t in range(2):
            assert torch.allclose(var[t], ref[0][t], rtol=1e-5, atol=1e-6), f"round {rnd} var {t}"
            assert torch.allclose(m[t], ref[1][t], rtol=1e-5, atol=1e-7)
            assert torch.allclose(v[t], ref[2][t], rtol=1e-5, atol=1e-7)
        assert torch.equal(shadow.float().view(-1), var[0].to(torch.bfloat16).float())
        assert abs(ctrl.view(torch.float32)[2].item() - b1p) < 1e-6
        produced, consumed = flags[:W].tolist(), flags[32:32 + W].tolist()
        assert produced == consumed == [int(s.item()) for s in seq]
    assert [int(s.item()) for s in seq] == [2, 2, 2]
    # a worker two pushes ahead: the drain retires ONE buffer per call, oldest first
    ga, gb = [torch.randn(n0, device=dev), torch.randn(n1, device=dev)], \
             [torch.randn(n0, device=dev), torch.randn(n1, device=dev)]
    push(1, ga)
    push(1, gb)
    assert int(minc.item()) == 2                    # what the owner had consumed when push 4 was published
    for g in (ga, gb):
        drain()
        for t in range(2):
            ref[0][t], ref[1][t], ref[2][t] = _adam_ref(ref[0][t], ref[1][t], ref[2][t], g[t],
                                                        lr, b1, b2, eps, b1p, b2p)
        b1p, b2p = b1p * b1, b2p * b2
        assert torch.allclose(var[0], ref[0][0], rtol=1e-5, atol=1e-6)
    assert flags[32 + 1].item() == 4 and ctrl.view(torch.int64)[0].item() == total + 2


@pytest.mark.parametrize("cols,ld", [(64, 64), (10, 16), (40, 40)])
def test_mailbox_drain_broadcasts_into_worker_mirrors(K, cols, ld):
    """Owner pushes, worker never pulls: the drain stores every updated element into each registered
    mirror (bf16 row-padded and fp32 forms, at the slice's place inside the whole variable) and then
    the header words the worker would otherwise read over NVLink. A paused shard and a held lock
    make the drain a no-op."""
    from distributedtensorflow_b200.ops import _lib as L
    torch.manual_seed(5)
    dev = "cuda"
    W, depth, shard = 2, 2, 3
    rows_total, row0, rows = 96, 32, 40            # this shard owns rows [32, 72) of a [96, cols] matrix
    n0, nb = rows * cols, 24                       # ... and elements [8, 32) of a 40-element vector
    off1 = (n0 * 4 + 255) // 256 * 256
    stride = off1 + 256
    slots = torch.zeros(W * depth * stride // 4, device=dev)
    flags = torch.zeros(128, dtype=torch.int32, device=dev)
    todo = torch.zeros(32, dtype=torch.int32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    ctrl = torch.zeros(64, dtype=torch.int32, device=dev)
    lr = 0.5
    var = [torch.randn(n0, device=dev), torch.randn(nb, device=dev)]
    shadow = torch.zeros(rows, ld, dtype=torch.bfloat16, device=dev)
    # two mirrors: header (512 B) | bf16 [rows_total, ld] | fp32 [40]
    bf_off, f32_off = 512, 512 + (rows_total * ld * 2 + 255) // 256 * 256
    mbytes = f32_off + 256
    mirrors = [torch.zeros(mbytes // 4, dtype=torch.int32, device=dev) for _ in range(W)]
    fl64 = flags.view(torch.int64)
    fl64[48 + 0] = mirrors[0].data_ptr()           # worker 0 and worker 1 registered
    fl64[48 + 1] = mirrors[1].data_ptr()
    items = K.make_apply_items([
        dict(var=var[0].data_ptr(), shadow=shadow.data_ptr(), grad=slots.data_ptr(), n=n0, cols=cols, shadow_ld=ld),
        dict(var=var[1].data_ptr(), shadow=None, grad=slots.data_ptr() + off1, n=nb, cols=nb, shadow_ld=nb)])
    mitems = K.make_mirror_items([(bf_off, -1, row0 * cols), (-1, f32_off, 8)])
    hyper = K.make_hyper(lr)
    seq = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(W)]
    pcount = torch.zeros(4, dtype=torch.int32, device=dev)
    minc = torch.zeros(1, dtype=torch.int32, device=dev)

    def push(w, g0, g1):
        base = slots.data_ptr() + w * depth * stride
        g1p = torch.zeros(32, device=dev)
        g1p[:nb] = g1
        K.mailbox_push([(g0.data_ptr(), base, n0 * 4, stride)], depth, w, [flags.data_ptr()], False,
                       seq[w].data_ptr(), None, None)
        K.mailbox_push([(g1p.data_ptr(), base + off1, 32 * 4, stride)], depth, w, [flags.data_ptr()], True,
                       seq[w].data_ptr(), pcount.data_ptr(), minc.data_ptr())
        torch.cuda.synchronize()

    def drain():
        K.mailbox_drain(L.OPT_SGD, items, 2, stride, depth, W, hyper, None, ctrl.data_ptr(),
                        flags.data_ptr(), todo.data_ptr(), counter.data_ptr(), None,
                        mirror_items=mitems, shard_index=shard, max_ctas=2 if cols == 64 else 0)
        torch.cuda.synchronize()

    ref = [x.clone() for x in var]
    g = [(torch.randn(n0, device=dev), torch.randn(nb, device=dev)) for _ in range(3)]
    push(0, *g[0])
    push(1, *g[1])
    # paused shard: nothing is claimed, nothing applied
    flags[L.MBOX_OFF_PAUSED // 4] = 1
    drain()
    assert torch.equal(var[0], ref[0]) and flags[32].item() == 0 and todo[16].item() == 0
    flags[L.MBOX_OFF_PAUSED // 4] = 0
    # a drain in flight elsewhere holds the lock: this launch backs off
    flags[L.MBOX_OFF_LOCK // 4] = 1
    drain()
    assert torch.equal(var[0], ref[0]) and flags[32].item() == 0
    flags[L.MBOX_OFF_LOCK // 4] = 0
    drain()
    for k in (0, 1):
        ref[0] -= lr * g[k][0]
        ref[1] -= lr * g[k][1]
    assert torch.allclose(var[0], ref[0], rtol=1e-6, atol=1e-6) and torch.allclose(var[1], ref[1], rtol=1e-6, atol=1e-6)
    assert flags[L.MBOX_OFF_LOCK // 4].item() == 0 and flags[L.MBOX_OFF_APPLIED // 4].item() == 2

    def check_mirrors(applied, consumed):
        want_bf = var[0].view(rows, cols).to(torch.bfloat16)
        for w, mir in enumerate(mirrors):
            raw = mir.view(torch.uint8)
            bf = raw[bf_off:bf_off + rows_total * ld * 2].view(torch.bfloat16).view(rows_total, ld)
            assert torch.equal(bf[row0:row0 + rows, :cols], want_bf), f"mirror {w} bf16"
            assert bf[:row0].abs().sum().item() == 0 and bf[row0 + rows:].abs().sum().item() == 0
            if ld > cols:
                assert bf[:, cols:].abs().sum().item() == 0            # the row padding is never written
            f32 = raw[f32_off:f32_off + 160].view(torch.float32)
            assert torch.equal(f32[8:32], var[1]) and f32[:8].abs().sum().item() == 0 and f32[32:].abs().sum().item() == 0
            hdr = mir[:128]
            assert hdr.view(torch.int64)[0].item() == applied                      # global_step
            assert hdr[L.MIRROR_OFF_VERSION // 4 + shard].item() == applied
            assert hdr[L.MIRROR_OFF_CONSUMED // 4 + shard].item() == consumed[w]
        assert torch.equal(shadow[:, :cols], want_bf)

    check_mirrors(2, [1, 1])
    # worker 1 unregisters: its mirror is left alone from now on
    fl64[48 + 1] = 0
    keep = mirrors[1].clone()
    push(0, *g[2])
    drain()
    ref[0] -= lr * g[2][0]
    assert torch.allclose(var[0], ref[0], rtol=1e-6, atol=1e-6)
    assert torch.equal(mirrors[1], keep)
    mirrors_all, mirrors[:] = list(mirrors), mirrors[:1]
    check_mirrors(3, [2])
    # the push kernel takes flow control and global_step from the worker's own header
    hdr_ptr = mirrors_all[0].data_ptr()
    step_out = torch.zeros(1, dtype=torch.int64, device=dev)
    base = slots.data_ptr()
    K.mailbox_push([(g[0][0].data_ptr(), base, n0 * 4, stride)], depth, 0, [flags.data_ptr()] * (shard + 1), True,
                   seq[0].data_ptr(), pcount.data_ptr(), minc.data_ptr(), ctrl_step=None,
                   step_out=step_out.data_ptr(), my_hdr=hdr_ptr)
    torch.cuda.synchronize()
    assert int(step_out.item()) == 3
    assert int(minc.item()) == 0          # min over header consumed[0..3]: shards 0..2 never wrote theirs


def test_mailbox_push_device_flow_control_waits_for_the_owner(K):
    """Bounded staleness on the device: a push that would overwrite a buffer its owner has not
    consumed yet parks (polling consumed[] in the worker's own mirror header) until the owner's
    drain moves that word - or until the worker's host raises the abort word."""
    import time
    from distributedtensorflow_b200.ops import _lib as L
    dev = "cuda"
    depth, owner, n = 2, 5, 4096
    hdr = torch.zeros(128, dtype=torch.int32, device=dev)
    flags = torch.zeros(128, dtype=torch.int32, device=dev)
    slots = torch.zeros(depth * n, device=dev)
    seq = torch.zeros(1, dtype=torch.int32, device=dev)
    pc = torch.zeros(4, dtype=torch.int32, device=dev)
    minc = torch.zeros(1, dtype=torch.int32, device=dev)
    grads = [torch.full((n,), float(k + 1), device=dev) for k in range(4)]
    side, poke = torch.cuda.Stream(), torch.cuda.Stream()

    def push(k):
        with torch.cuda.stream(side):
            K.mailbox_push([(grads[k].data_ptr(), slots.data_ptr(), n * 4, n * 4, 0, owner)], depth, 0,
                           [flags.data_ptr()], True, seq.data_ptr(), pc.data_ptr(), minc.data_ptr(),
                           my_hdr=hdr.data_ptr())

    push(0)
    push(1)                       # buffers 0 and 1: nothing to wait for
    side.synchronize()
    assert slots[:n].eq(1).all() and slots[n:].eq(2).all() and seq.item() == 2
    push(2)                       # buffer 0 again: needs consumed[owner] >= 1
    time.sleep(0.05)
    assert not side.query() and bool((slots[:n].cpu() == 1).all())     # (read back with a copy, not a kernel)
    # the words are moved by a COPY (copy engine), not by a kernel: a kernel queued behind the parked
    # one could wait for it for ever, a copy cannot
    one = torch.ones(1, dtype=torch.int32).pin_memory()
    with torch.cuda.stream(poke):
        hdr[L.MIRROR_OFF_CONSUMED // 4 + owner: L.MIRROR_OFF_CONSUMED // 4 + owner + 1].copy_(one, non_blocking=True)
    poke.synchronize()
    side.synchronize()
    assert slots[:n].eq(3).all() and seq.item() == 3
    push(3)                       # buffer 1: needs consumed >= 2; the host gives up instead
    time.sleep(0.05)
    assert not side.query()
    with torch.cuda.stream(poke):
        hdr[L.MIRROR_OFF_ABORT // 4: L.MIRROR_OFF_ABORT // 4 + 1].copy_(one, non_blocking=True)
    poke.synchronize()
    side.synchronize()
    assert seq.item() == 4


@pytest.mark.parametrize("B,H,dense", [(128, 1024, False), (100, 100, True), (37, 300, False)])
def test_head_fused_matches_separate_ops(K, B, H, dense):
    """logits + softmax-xent + dlogits + bias grads + masked dX of the last layer in one kernel,
    against a plain PyTorch fp32 computation on the same bf16-rounded operands."""
    torch.manual_seed(B + H)
    C, dev = 10, "cuda"
    ldh = (H + 7) // 8 * 8
    act = torch.zeros(B, ldh, dtype=torch.bfloat16, device=dev)
    act[:, :H] = torch.relu(torch.randn(B, H, device=dev)).to(torch.bfloat16)
    W = torch.zeros(H, 16, dtype=torch.bfloat16, device=dev)
    W[:, :C] = (torch.randn(H, C, device=dev) * 0.1).to(torch.bfloat16)
    bias = torch.randn(C, device=dev) * 0.1
    if dense:
        lab = torch.randint(0, C, (B,), device=dev)
        labels = torch.nn.functional.one_hot(lab, C).float()
    else:
        labels = torch.randint(0, C, (B,), device=dev, dtype=torch.int32)
    logits = torch.zeros(B, C, device=dev)
    loss = torch.zeros(1, device=dev)
    dlogits = torch.zeros(B, 16, dtype=torch.bfloat16, device=dev)
    db = torch.zeros(C, device=dev)
    dact = torch.zeros(B, ldh, dtype=torch.bfloat16, device=dev)
    dbp = torch.zeros(H, device=dev)
    dw = torch.zeros(H, C, device=dev)
    K.head_fused(act, H, W.data_ptr(), 16, bias.data_ptr(), labels, B, C, logits, loss,
                 dlogits=dlogits, dbias_accum=db, dact=dact, dprev_bias_accum=dbp, dw_accum=dw)
    torch.cuda.synchronize()
    a32, w32 = act[:, :H].float(), W[:, :C].float()
    z = (a32 @ w32 + bias).requires_grad_(True)
    if dense:
        p = torch.softmax(z, 1)
        ref_loss = -(labels * torch.log(torch.clamp(p, 1e-10, 1.0))).sum()
    else:
        ref_loss = torch.nn.functional.cross_entropy(z, labels.long())
    ref_loss.backward()
    g = z.grad
    assert torch.allclose(logits, z.detach(), rtol=1e-4, atol=1e-4)
    assert abs(loss.item() - ref_loss.item()) < 1e-3 * max(1.0, abs(ref_loss.item()))
    assert torch.allclose(dlogits[:, :C].float(), g, rtol=2e-2, atol=1e-6 + 4e-3 * g.abs().max().item())
    assert (dlogits[:, C:] == 0).all()
    assert torch.allclose(db, g.sum(0), rtol=1e-3, atol=1e-5 + 1e-3 * g.abs().max().item())
    assert _rel_err(dw, a32.t() @ dlogits[:, :C].float()) < 1e-3
    ref_dact = (dlogits[:, :C].float() @ w32.t()) * (a32 > 0)
    assert _rel_err(dact[:, :H].float(), ref_dact) < 1e-2
    assert (dact[:, H:] == 0).all()
    assert torch.allclose(dbp, dact[:, :H].float().sum(0), rtol=1e-3, atol=1e-5 + 1e-3 * ref_dact.abs().max().item())
    # inference form: logits + loss only
    logits2, loss2 = torch.zeros_like(logits), torch.zeros(1, device=dev)
    K.head_fused(act, H, W.data_ptr(), 16, bias.data_ptr(), labels, B, C, logits2, loss2)
    torch.cuda.synchronize()
    assert torch.equal(logits2, logits) and abs(loss2.item() - loss.item()) < 1e-5 * max(1.0, abs(loss.item()))


@pytest.mark.parametrize("M,N,K_,rps,block_n", [(512, 256, 128, 192, 0), (300, 10, 96, 300, 0), (1024, 1024, 256, 512, 256)])
def test_gemm_scatter_epilogue_stores_rows_into_mailbox_buffers(K, M, N, K_, rps, block_n):
    """dW GEMM whose epilogue is the gradient push: row m lands in slice m // rps, inside buffer
    (*seq % depth) of that slice's mailbox; the other buffers stay untouched."""
    torch.manual_seed(M + N)
    dev = "cuda"
    A = torch.randn(K_, M, device=dev)          # a_mn: stored [K, M]
    B = torch.randn(K_, N, device=dev)          # b_mn: stored [K, N]
    a, b = _padded(K_, M, fill=A), _padded(K_, N, fill=B)
    nsl = -(-M // rps)
    depth = 3
    stride = (rps * N * 4 + 255) // 256 * 256
    boxes = [torch.full((depth * stride // 4,), float("nan"), device=dev) for _ in range(nsl)]
    seq = torch.tensor([7], dtype=torch.int32, device=dev)       # 7 % 3 -> buffer 1
    K.gemm(a, b, None, M, N, K_, a_mn=True, b_mn=True, block_n=block_n,
           scatter=dict(rows_per_slice=rps, depth=depth, seq=seq.data_ptr(),
                        slices=[(bx.data_ptr(), stride) for bx in boxes]))
    torch.cuda.synchronize()
    ref = _bf16(A).float().t() @ _bf16(B).float()
    for i, bx in enumerate(boxes):
        rows = min(rps, M - i * rps)
        bufs = bx.view(depth, stride // 4)
        got = bufs[1, :rows * N].view(rows, N)
        assert _rel_err(got, ref[i * rps:i * rps + rows]) < 2e-3
        assert torch.isnan(bufs[0]).all() and torch.isnan(bufs[2]).all()
        assert torch.isnan(bufs[1, rows * N:]).all()
