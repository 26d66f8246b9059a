"""Kernel-level parity on the GPU: every C-ABI kernel against a plain PyTorch restatement of the same op.

Integer / index / byte work (preprocess, gathers) is compared bit-exactly; floating-point kernels with the
tolerance written next to each assert.  The step-level parity against the CPU oracle lives in
test_step_parity_gpu.py.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from b200step import ops as _ops
    from b200step.lib import get_lib

    get_lib()  # raises if libb200step.so is missing: no fallback
    return _ops


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).cuda()


# ------------------------------------------------------------------------------------------------ GEMMs
GEMM_SHAPES = [
    # (M, N, K)   M = batch rows, N = out features, K = in features
    (128, 256, 64),
    (256, 512, 448),
    (1000, 512, 512),     # ragged M
    (4096, 1000, 512),    # ragged N (output layer: 1000 classes)
    (333, 1000, 448),
    (4096, 512, 448),
    # T1 (`tabular_mlp_wide`, BASELINE config 4): 16 and 27 k-blocks against the 6-stage operand ring
    (4096, 1024, 1728),
    (4096, 1024, 1024),
    (4096, 1000, 1024),
    (333, 1024, 1728),
]


def _tols(dtype):
    # bf16: inputs exact in bf16, fp32 accumulate, output rounded to bf16 (rel 2^-8); tf32: 10-bit mantissa inputs
    return (2e-2, 2e-2) if dtype == torch.bfloat16 else (5e-3, 5e-3)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_linear_fwd(ops, dtype, M, N, K):
    x = _rand((M, K), dtype, 1)
    w = _rand((N, K), dtype, 2, 0.05)
    b = _rand((N,), torch.float32, 3)
    y = torch.full((M, N), float("nan"), dtype=dtype, device="cuda")
    for relu in (False, True):
        ops.linear_fwd(x, w, b, y, M=M, N=N, K=K, relu=relu)
        torch.cuda.synchronize()
        ref = x.double() @ w.double().t() + b.double()
        if relu:
            ref = ref.clamp_min(0)
        rtol, atol = _tols(dtype)
        torch.testing.assert_close(y.double(), ref, rtol=rtol, atol=atol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_linear_dgrad(ops, dtype, M, N, K):
    dy = _rand((M, N), dtype, 4)
    w = _rand((N, K), dtype, 5, 0.05)
    act = _rand((M, K), dtype, 6)
    dx = torch.full((M, K), float("nan"), dtype=dtype, device="cuda")
    rtol, atol = _tols(dtype)
    ops.linear_dgrad(dy, w, None, dx, M=M, N=N, K=K, relu=False)
    torch.cuda.synchronize()
    ref = dy.double() @ w.double()
    torch.testing.assert_close(dx.double(), ref, rtol=rtol, atol=atol * 2)
    ops.linear_dgrad(dy, w, act, dx, M=M, N=N, K=K, relu=True, dscale=2.0)
    torch.cuda.synchronize()
    ref = ref * (act.double() > 0) * 2.0
    torch.testing.assert_close(dx.double(), ref, rtol=rtol, atol=atol * 4)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES + [(65536, 512, 448)])
def test_linear_wgrad(ops, dtype, M, N, K):
    dy = _rand((M, N), dtype, 7, 0.1)
    x = _rand((M, K), dtype, 8)
    dw = torch.full((N, K), float("nan"), dtype=torch.float32, device="cuda")
    ops.linear_wgrad(dy, x, dw, M=M, N=N, K=K)
    torch.cuda.synchronize()
    ref = dy.double().t() @ x.double()
    # fp32 result of an M-term sum: error grows ~ sqrt(M) * eps_in * |dy||x|
    scale = (M ** 0.5) * 0.1
    tol = (1e-5 if dtype == torch.bfloat16 else 2e-3) * scale + 1e-5
    assert (dw.double() - ref).abs().max().item() <= tol * 4
    # accumulate flag adds on top of what is there
    ops.linear_wgrad(dy, x, dw, M=M, N=N, K=K, accumulate=True)
    torch.cuda.synchronize()
    assert (dw.double() - 2 * ref).abs().max().item() <= tol * 8


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,K", [(1000, 512, 448), (4096, 1000, 512), (65536, 512, 512), (4096, 1024, 1728)])
def test_relu_bitmask_and_fused_bias_grad(ops, dtype, M, N, K):
    """fwd writes 1 bit per output (y > 0); the next layer's dgrad consumes it instead of re-reading the
    activation, and emits the column sums of what it stores (= bias gradient of the layer below)."""
    x = _rand((M, K), dtype, 1)
    w = _rand((N, K), dtype, 2, 0.05)
    y = torch.empty((M, N), dtype=dtype, device="cuda")
    words = ((N + 31) // 32 + 3) // 4 * 4
    mask = torch.zeros((M, words), dtype=torch.int32, device="cuda")
    ops.linear_fwd(x, w, None, y, M=M, N=N, K=K, relu=True, relu_mask=mask)
    torch.cuda.synchronize()
    bits = ((mask[:, :, None] >> torch.arange(32, device="cuda", dtype=torch.int32)) & 1).reshape(M, -1)[:, :N]
    assert torch.equal(bits.bool(), y > 0)                       # bit-exact
    # backward of a layer whose INPUT was y: dx[M,N] = (dy2[M,N2] @ w2[N2,N]) * (y > 0)
    N2 = 256
    dy2 = _rand((M, N2), dtype, 4)
    w2 = _rand((N2, N), dtype, 5, 0.05)
    dx_act = torch.empty((M, N), dtype=dtype, device="cuda")
    dx_mask = torch.full((M, N), float("nan"), dtype=dtype, device="cuda")
    db = torch.full((N,), float("nan"), device="cuda")
    scratch = torch.empty(ops.dgrad_scratch_floats(M, N), device="cuda")
    ops.linear_dgrad(dy2, w2, y, dx_act, M=M, N=N2, K=N, relu=True)
    ops.linear_dgrad(dy2, w2, None, dx_mask, M=M, N=N2, K=N, relu=True, relu_mask=mask, dbias=db, scratch=scratch)
    torch.cuda.synchronize()
    assert torch.equal(dx_act, dx_mask)                          # same kernel math, mask from bits == mask from y
    torch.testing.assert_close(db.double(), dx_mask.double().sum(0), rtol=1e-4, atol=1e-3 * (M ** 0.5) * 1e-2)
    db2 = torch.empty_like(db)
    ops.linear_dgrad(dy2, w2, None, dx_mask, M=M, N=N2, K=N, relu=True, relu_mask=mask, dbias=db2, scratch=scratch)
    torch.cuda.synchronize()
    assert torch.equal(db, db2)                                  # deterministic reduction order
    # without scratch: float atomics onto a zeroed vector (what the training step uses)
    db3 = torch.zeros_like(db)
    ops.linear_dgrad(dy2, w2, None, dx_mask, M=M, N=N2, K=N, relu=True, relu_mask=mask, dbias=db3)
    torch.cuda.synchronize()
    torch.testing.assert_close(db3, db, rtol=1e-4, atol=1e-3 * (M ** 0.5) * 1e-2)


T0_LAYERS = [(1000, 512), (512, 512), (512, 448)]                                        # (N, K), output layer first
T1_LAYERS = [(1000, 1024), (1024, 1024), (1024, 1024), (1024, 1024), (1024, 1728)]     # five problems in one launch


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,shapes", [(4096, T0_LAYERS), (65536, T0_LAYERS), (1000, T0_LAYERS), (4096, T1_LAYERS),
                                      (333, T1_LAYERS)], ids=["t0-4096", "t0-65536", "t0-1000", "t1-4096", "t1-333"])
def test_linear_wgrad_grouped(ops, dtype, M, shapes):
    """All layers' weight gradients in one launch == the per-layer kernel (same tiles, same arithmetic)."""
    dys = [_rand((M, n), dtype, 70 + i, 0.1) for i, (n, k) in enumerate(shapes)]
    xs = [_rand((M, k), dtype, 80 + i) for i, (n, k) in enumerate(shapes)]
    dws = [torch.zeros((n, k), dtype=torch.float32, device="cuda") for (n, k) in shapes]
    ops.linear_wgrad_grouped(dys, xs, dws, M=M, Ns=[n for n, _ in shapes], Ks=[k for _, k in shapes])
    torch.cuda.synchronize()
    for dy, x, dw, (n, k) in zip(dys, xs, dws, shapes):
        ref = dy.double().t() @ x.double()
        scale = (M ** 0.5) * 0.1
        tol = (1e-5 if dtype == torch.bfloat16 else 2e-3) * scale + 1e-5
        assert (dw.double() - ref).abs().max().item() <= tol * 4
    # accumulates onto what is there
    ops.linear_wgrad_grouped(dys, xs, dws, M=M, Ns=[n for n, _ in shapes], Ks=[k for _, k in shapes])
    torch.cuda.synchronize()
    ref = dys[0].double().t() @ xs[0].double()
    assert (dws[0].double() - 2 * ref).abs().max().item() <= tol * 8 * 10


def test_gemm_rejects_misaligned(ops):
    from b200step.lib import B200Error

    x = torch.zeros(128, 60, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(64, 60, dtype=torch.bfloat16, device="cuda")
    y = torch.zeros(128, 64, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(B200Error):
        ops.linear_fwd(x, w, None, y, M=128, N=64, K=60, relu=False)


def test_dropout_mask_matches_philox_oracle(ops):
    from oracle.philox import dropout_keep_mask

    M, N, K = 256, 512, 64
    x = _rand((M, K), torch.bfloat16, 11)
    w = _rand((N, K), torch.bfloat16, 12, 0.1)
    y0 = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    y1 = torch.empty_like(y0)
    step = torch.tensor([3], dtype=torch.int32, device="cuda")
    ops.linear_fwd(x, w, None, y0, M=M, N=N, K=K, relu=False)
    ops.linear_fwd(x, w, None, y1, M=M, N=N, K=K, relu=False, dropout_p=0.25, seed=1234, step=step, layer_id=2)
    torch.cuda.synchronize()
    keep = torch.from_numpy(dropout_keep_mask(M, N, p=0.25, seed=1234, step=3, layer_id=2)).cuda()
    ref = torch.where(keep, (y0.float() / (1 - 0.25)), torch.zeros((), device="cuda")).to(torch.bfloat16)
    # the kernel scales the fp32 accumulator before rounding; y0 is already rounded -> 1 bf16 ulp slack
    torch.testing.assert_close(y1.float(), ref.float(), rtol=1.6e-2, atol=1e-3)
    assert torch.equal(y1 != 0, keep & (y0 != 0)) or ((y1 != 0) ^ (keep & (y0 != 0))).sum().item() <= 2


# ------------------------------------------------------------------------------------------- preprocess
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("batch,nd,nc,rows,dim,ldx", [
    (4096, 13, 26, 1000, 16, 448),
    (1000, 13, 26, 1000, 16, 448),   # ragged batch (last CTA tile partial)
    (257, 13, 26, 65536, 64, 1728),  # T1 row shape
    (64, 13, 0, 1, 4, 16),           # no categorical features
    (96, 5, 40, 50, 8, 328),         # > 32 tables
])
def test_preprocess_bit_exact(ops, dtype, batch, nd, nc, rows, dim, ldx):
    g = torch.Generator().manual_seed(5)
    nrows = 3 * batch + 17
    dense = (torch.randn(nrows, nd, generator=g) * 3 + 1).cuda()
    cat = torch.randint(0, rows, (nrows, max(nc, 1)), generator=g)[:, :nc].contiguous().cuda()
    label = torch.randint(0, 10, (nrows,), generator=g).cuda()
    mean = dense.mean(0).contiguous()
    std = dense.std(0).contiguous()
    tables = torch.randn(max(nc, 1) * rows * dim, generator=g).cuda()
    steps = 2
    row_idx = torch.randperm(nrows, generator=g)[: steps * batch].cuda()
    cursor = torch.tensor([1], dtype=torch.int32, device="cuda")
    x = torch.full((batch, ldx), float("nan"), dtype=dtype, device="cuda")
    y = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
    idx = torch.full((batch, max(nc, 1)), -1, dtype=torch.int32, device="cuda")
    ops.preprocess(dense, cat, label, row_idx, cursor, mean, std, tables, x, y, idx, batch=batch, num_dense=nd,
                   num_cat=nc, table_rows=rows, embed_dim=dim)
    torch.cuda.synchronize()
    r = row_idx[batch: 2 * batch]
    ref = torch.zeros(batch, ldx, dtype=torch.float32, device="cuda")
    # row layout: [embeddings | standardised dense | zero pad]
    ref[:, nc * dim: nc * dim + nd] = dense[r].sub(mean).div(std)  # reference rule: tensor.sub_(mean).div_(std)
    if nc:
        tv = tables.view(nc, rows, dim)
        emb = torch.stack([tv[t][cat[r, t]] for t in range(nc)], 1).reshape(batch, nc * dim)
        ref[:, : nc * dim] = emb
        assert torch.equal(idx[:, :nc].long(), cat[r])
    ref = ref.to(dtype)
    assert torch.equal(x.view(torch.int16 if dtype == torch.bfloat16 else torch.int32),
                       ref.view(torch.int16 if dtype == torch.bfloat16 else torch.int32))
    assert torch.equal(y, label[r])


@pytest.mark.parametrize("budget", [116, 8, 2])
def test_sm_budget_changes_grids_not_results(ops, budget):
    """b200_set_sm_budget: K-pre switches to one fat CTA per budgeted SM, the GEMMs to fewer CTA pairs; outputs are
    bit-identical to the full-device launch (wgrad: same split-K sums up to float atomics order)."""
    g = torch.Generator().manual_seed(11)
    batch, nd, nc, rows, dim, ldx = 5000, 13, 26, 1000, 16, 448
    dense = (torch.randn(batch, nd, generator=g) * 3 + 1).cuda()
    cat = torch.randint(0, rows, (batch, nc), generator=g).cuda()
    label = torch.randint(0, 10, (batch,), generator=g).cuda()
    mean, std = dense.mean(0).contiguous(), dense.std(0).contiguous()
    tables = torch.randn(nc * rows * dim, generator=g).cuda()
    row_idx = torch.randperm(batch, generator=g).cuda()
    w = _rand((512, ldx), torch.bfloat16, 3, 0.05)
    bias = _rand((512,), torch.float32, 4, 0.1)

    def run():
        x = torch.full((batch, ldx), float("nan"), dtype=torch.bfloat16, device="cuda")
        y = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        idx = torch.full((batch, nc), -1, dtype=torch.int32, device="cuda")
        ops.preprocess(dense, cat, label, row_idx, None, mean, std, tables, x, y, idx, batch=batch, num_dense=nd,
                       num_cat=nc, table_rows=rows, embed_dim=dim)
        h = torch.empty(batch, 512, dtype=torch.bfloat16, device="cuda")
        ops.linear_fwd(x, w, bias, h, M=batch, N=512, K=ldx, relu=True)
        dw = torch.zeros(512, ldx, dtype=torch.float32, device="cuda")
        ops.linear_wgrad(h, x, dw, M=batch, N=512, K=ldx, accumulate=True)
        torch.cuda.synchronize()
        return x, y, idx, h, dw

    full = run()
    ops.set_sm_budget(budget)
    try:
        part = run()
    finally:
        ops.set_sm_budget(0)
    for a, b in zip(full[:4], part[:4]):
        assert torch.equal(a.view(torch.int16) if a.dtype == torch.bfloat16 else a,
                           b.view(torch.int16) if b.dtype == torch.bfloat16 else b)
    assert (full[4] - part[4]).abs().max().item() <= 1e-3 * full[4].abs().max().item()
    again = run()                       # budget cleared: back to the default grids
    assert torch.equal(again[0].view(torch.int16), full[0].view(torch.int16))


# ------------------------------------------------------------------------------------------------- loss
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("batch,C", [(4096, 1000), (1000, 1000), (37, 16), (65536, 1000), (512, 1024)])
def test_ce_loss_grad(ops, dtype, batch, C):
    logits = _rand((batch, C), dtype, 21, 2.0)
    labels = torch.randint(0, min(C, 10), (batch,), generator=torch.Generator().manual_seed(22)).cuda()
    dlogits = torch.full((batch, C), float("nan"), dtype=dtype, device="cuda")
    loss = torch.zeros(1, device="cuda")
    dbias = torch.full((C,), float("nan"), device="cuda")
    scratch = torch.empty(ops.ce_scratch_floats(C), device="cuda")
    scale = 1.0 / (batch * 2)  # world = 2
    ops.ce_loss_grad(logits, labels, dlogits, loss, dbias, scratch, batch=batch, num_classes=C, grad_scale=scale)
    torch.cuda.synchronize()
    z = logits.double().requires_grad_(True)
    ref_loss = torch.nn.functional.cross_entropy(z, labels)
    ref_loss.backward()
    ref_d = z.grad * (batch * scale)
    # loss: fp32 exp/log + fp32 tree sums -> rel 1e-5
    assert abs(loss.item() - ref_loss.item()) <= 1e-5 * max(1.0, abs(ref_loss.item()))
    if dtype == torch.float32:
        torch.testing.assert_close(dlogits.double(), ref_d, rtol=1e-4, atol=1e-9)
    else:  # one bf16 rounding of the fp32 gradient
        torch.testing.assert_close(dlogits.double(), ref_d, rtol=8e-3, atol=1e-9)
    # bias gradient = column sums of what was stored
    torch.testing.assert_close(dbias.double(), dlogits.double().sum(0), rtol=1e-4, atol=1e-7)
    # deterministic: bit-identical on a second call
    loss2 = torch.zeros(1, device="cuda")
    dbias2 = torch.empty_like(dbias)
    ops.ce_loss_grad(logits, labels, dlogits, loss2, dbias2, scratch, batch=batch, num_classes=C, grad_scale=scale)
    torch.cuda.synchronize()
    assert torch.equal(loss, loss2) and torch.equal(dbias, dbias2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("M,N", [(4096, 512), (1000, 1000), (65536, 512), (7, 8), (300, 4096)])
def test_colsum(ops, dtype, M, N):
    dy = _rand((M, N), dtype, 31)
    out = torch.full((N,), float("nan"), device="cuda")
    scratch = torch.empty(ops.colsum_scratch_floats(N), device="cuda")
    ops.colsum(dy, out, scratch, M=M, N=N)
    torch.cuda.synchronize()
    ref = dy.double().sum(0)
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-3 * (M ** 0.5) * 1e-2)


# -------------------------------------------------------------------------------------------- embedding
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("batch,nc,rows,dim,skew", [
    (4096, 26, 1000, 16, False),     # tables fit shared memory: privatised kernel
    (65536, 26, 1000, 16, False),    # BASELINE config 3's batch
    (65536, 26, 1000, 16, True),     # Zipf-like ids: a few hot rows take most of the adds
    (5000, 26, 17, 8, False),        # tiny tables, ragged batch slices
    (1000, 26, 1000, 16, False),     # small batch: plain global scatter
    (4096, 26, 4096, 64, False),     # T1 row shape, 1 MB per table: plain global scatter
], ids=["b4096", "b65536", "b65536-skew", "tiny-tables", "b1000-global", "t1-rows-global"])
def test_emb_bwd(ops, dtype, batch, nc, rows, dim, skew):
    col0 = 16  # embedding columns may start at any multiple of 4 (0 in the training step)
    ldx = (col0 + nc * dim + 13 + 63) // 64 * 64
    dx = _rand((batch, ldx), dtype, 41, 0.01)
    g = torch.Generator().manual_seed(42)
    if skew:
        probs = torch.arange(1, rows + 1, dtype=torch.float64).pow(-1.05)
        idx = torch.multinomial(probs / probs.sum(), batch * nc, replacement=True, generator=g).view(batch, nc)
    else:
        idx = torch.randint(0, rows, (batch, nc), generator=g)
    idx = idx.int().cuda()
    dt = torch.zeros(nc * rows * dim, device="cuda")
    ops.emb_bwd(dx, idx, dt, batch=batch, emb_col0=col0, num_cat=nc, table_rows=rows, embed_dim=dim)
    torch.cuda.synchronize()
    ref = torch.zeros(nc, rows, dim, dtype=torch.float64, device="cuda")
    for t in range(nc):
        ref[t].index_add_(0, idx[:, t].long(), dx[:, col0 + t * dim: col0 + (t + 1) * dim].double())
    # fp32 atomics in arbitrary order: |err| <= n_adds * eps * max|term|
    n_adds = max(1.0, batch / rows) * (20 if skew else 1)
    torch.testing.assert_close(dt.view(nc, rows, dim).double(), ref, rtol=1e-5, atol=1e-6 * max(1.0, n_adds / 4))
    # adds onto what is there (the step zeroes the table gradients in its optimizer pass, chunks accumulate)
    ops.emb_bwd(dx, idx, dt, batch=batch, emb_col0=col0, num_cat=nc, table_rows=rows, embed_dim=dim)
    torch.cuda.synchronize()
    torch.testing.assert_close(dt.view(nc, rows, dim).double(), 2 * ref, rtol=2e-5, atol=2e-6 * max(1.0, n_adds / 4))


# ------------------------------------------------------------------------------------------- optimizers
@pytest.mark.parametrize("n", [1024, 1411816 + 3, 17])
def test_sgd_matches_torch(ops, n):
    g = torch.Generator().manual_seed(51)
    p0 = torch.randn(n, generator=g)
    n_shadow = (n // 2) // 4 * 4
    p = p0.clone().cuda()
    buf = torch.zeros(n, device="cuda")
    shadow = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=1e-4)
    for it in range(3):
        grad = torch.randn(n, generator=g) * 0.1
        ref_p.grad = grad.clone()
        opt.step()
        ops.sgd_step(p, grad.cuda(), buf, shadow, n_shadow=n_shadow, lr=0.1, momentum=0.9, weight_decay=1e-4)
    torch.cuda.synchronize()
    # same fp32 operation order as torch's SGD; FMA contraction may differ by 1 ulp per op
    torch.testing.assert_close(p.cpu(), ref_p.data, rtol=1e-6, atol=1e-7)
    assert torch.equal(shadow[:n_shadow], p[:n_shadow].to(torch.bfloat16))
    assert torch.count_nonzero(shadow[n_shadow:]).item() == 0


@pytest.mark.parametrize("n", [4096, 100003])
def test_adam_matches_torch(ops, n):
    g = torch.Generator().manual_seed(61)
    p0 = torch.randn(n, generator=g)
    p = p0.clone().cuda()
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    step = torch.zeros(1, dtype=torch.int32, device="cuda")
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref_p], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4)
    for it in range(4):
        grad = torch.randn(n, generator=g) * 0.1
        ref_p.grad = grad.clone()
        opt.step()
        ops.adam_step(p, grad.cuda(), m, v, None, step, n_shadow=0, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
                      weight_decay=1e-4)
        ops.advance_step(None, step, 1)
    torch.cuda.synchronize()
    assert step.item() == 4
    torch.testing.assert_close(p.cpu(), ref_p.data, rtol=2e-6, atol=2e-7)


def test_cast_and_advance(ops):
    src = _rand((1000003,), torch.float32, 71)
    dst = torch.zeros(1000003, dtype=torch.bfloat16, device="cuda")
    ops.cast_bf16(src, dst)
    cursor = torch.tensor([2], dtype=torch.int32, device="cuda")
    step = torch.tensor([9], dtype=torch.int32, device="cuda")
    ops.advance_step(cursor, step, 3)
    torch.cuda.synchronize()
    assert torch.equal(dst, src.to(torch.bfloat16))
    assert cursor.item() == 0 and step.item() == 10


# ----------------------------------------------------------- inputs torch would reject / ignore_index semantics
def test_ce_ignore_index_and_bad_labels(ops):
    """nn.CrossEntropyLoss defaults (loss.py:1379-1387): targets equal to ignore_index=-100 add neither loss nor
    gradient and the mean runs over the remaining rows; any other target outside [0, C) makes torch raise - the
    kernel drops the row and sets B200_STATUS_BAD_LABEL."""
    import torch.nn.functional as F

    B, C = 777, 1000
    status = ops.status_word()
    status.zero_()
    for dtype in (torch.float32, torch.bfloat16):
        logits = _rand((B, C), dtype, 21, 3.0)
        g = torch.Generator().manual_seed(22)
        labels = torch.randint(0, C, (B,), generator=g)
        labels[torch.randperm(B, generator=g)[:100]] = -100
        labels = labels.cuda()
        n_valid = torch.tensor([int((labels != -100).sum())], dtype=torch.int32, device="cuda")
        dl = torch.full_like(logits, float("nan"))
        loss = torch.zeros(1, device="cuda")
        db = torch.zeros(C, device="cuda")
        scratch = torch.empty(ops.ce_scratch_floats(C), device="cuda")
        ops.ce_loss_grad(logits, labels, dl, loss, db, scratch, batch=B, num_classes=C, grad_scale=1.0 / B,
                         n_valid=n_valid)
        torch.cuda.synchronize()
        z = logits.double().requires_grad_(True)
        ref = F.cross_entropy(z, labels)            # ignore_index=-100, reduction="mean"
        ref.backward()
        assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
        tol = 1e-6 if dtype == torch.float32 else 2e-2
        torch.testing.assert_close(dl.double(), z.grad, rtol=tol, atol=tol / B)
        assert torch.count_nonzero(dl[labels == -100]).item() == 0
        assert int(status.item()) == 0
        # a label torch would raise on: dropped like an ignored row, and reported
        bad = labels.clone()
        bad[5] = C + 3
        ops.ce_loss_grad(logits, bad, dl, loss, db, scratch, batch=B, num_classes=C, grad_scale=1.0 / B)
        torch.cuda.synchronize()
        assert int(status.item()) == ops.STATUS_BAD_LABEL
        assert torch.count_nonzero(dl[5]).item() == 0 and torch.isfinite(dl).all()
        status.zero_()
    # no valid row at all: torch's mean is 0/0
    labels = torch.full((64,), -100, dtype=torch.int64, device="cuda")
    logits = _rand((64, C), torch.float32, 23)
    dl = torch.full_like(logits, float("nan"))
    ops.ce_loss_grad(logits, labels, dl, loss, None, scratch, batch=64, num_classes=C, grad_scale=1.0 / 64,
                     n_valid=torch.zeros(1, dtype=torch.int32, device="cuda"))
    torch.cuda.synchronize()
    assert torch.isnan(loss).all() and torch.count_nonzero(dl).item() == 0


@pytest.mark.parametrize("id_dtype", [torch.int64, torch.int32])
def test_preprocess_flags_out_of_range_ids_and_counts_labels(ops, id_dtype):
    """nn.Embedding raises IndexError on an id outside its table; K-pre clamps it to row 0 (so neither it nor the
    embedding backward, which reuses the stored ids, leaves the table) and sets B200_STATUS_BAD_CATEGORICAL_ID.
    label_stats[0] = rows whose label is not ignore_index, re-armed after every launch."""
    batch, nd, nc, rows, dim, ldx = 1000, 13, 26, 100, 16, 448
    g = torch.Generator().manual_seed(31)
    dense = torch.randn(batch, nd, generator=g).cuda()
    cat = torch.randint(0, rows, (batch, nc), generator=g).to(id_dtype)
    label = torch.randint(0, 10, (batch,), generator=g)
    label[::7] = -100
    mean, std = dense.mean(0).contiguous(), dense.std(0).contiguous()
    tables = torch.randn(nc * rows * dim, generator=g).cuda()
    row_idx = torch.arange(batch).cuda()
    status = ops.status_word()
    status.zero_()
    stats = torch.zeros(4, dtype=torch.int32, device="cuda")

    def run(cat_t):
        x = torch.full((batch, ldx), float("nan"), dtype=torch.bfloat16, device="cuda")
        y = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        idx = torch.full((batch, nc), -1, dtype=torch.int32, device="cuda")
        ops.preprocess(dense, cat_t.cuda(), label.cuda(), row_idx, None, mean, std, tables, x, y, idx, batch=batch,
                       num_dense=nd, num_cat=nc, table_rows=rows, embed_dim=dim, label_stats=stats)
        torch.cuda.synchronize()
        return x, idx

    for _ in range(2):                         # twice: the accumulator and ticket must re-arm themselves
        x, idx = run(cat)
        assert int(status.item()) == 0
        assert stats.tolist() == [int((label != -100).sum()), 0, 0, 0]
    bad = cat.clone()
    bad[3, 4] = rows            # one past the end
    bad[10, 0] = -1
    x, idx = run(bad)
    assert int(status.item()) == ops.STATUS_BAD_CATEGORICAL_ID
    assert idx[3, 4].item() == 0 and idx[10, 0].item() == 0 and torch.isfinite(x.float()).all()
    tv = tables.view(nc, rows, dim)
    assert torch.equal(x[3, 4 * dim: 5 * dim].float(), tv[4][0].to(torch.bfloat16).float())
    status.zero_()


# ------------------------------------------------------------------- output layer + cross-entropy in one kernel
LINEAR_CE_SHAPES = [
    # (M, N classes, K)
    (4096, 1000, 512),     # T0 output layer: a cluster of 4 CTA pairs, last tile ragged (1000 = 3*256 + 232)
    (65536, 1000, 512),    # BASELINE config 3's batch
    (333, 1000, 512),      # ragged rows
    (1000, 40, 64),        # one pair (the small test models)
    (1000, 512, 448),      # two pairs
    (512, 600, 128),       # three class tiles -> four pairs, the last one idle
    (777, 1024, 512),      # all 1024 classes
    (4096, 1000, 1024),    # T1 output layer (16 k-blocks)
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,K", LINEAR_CE_SHAPES)
def test_linear_ce_matches_linear_then_loss(ops, dtype, M, N, K):
    """b200_linear_ce (logits kept in tensor memory, fp32, never rounded) against b200_linear_fwd + b200_ce_loss_grad
    (logits through HBM in the storage dtype) and against torch in fp64."""
    import torch.nn.functional as F

    x = _rand((M, K), dtype, 41)
    w = _rand((N, K), dtype, 42, 0.08)
    b = _rand((N,), torch.float32, 43, 0.5)
    g = torch.Generator().manual_seed(44)
    labels = torch.randint(0, N, (M,), generator=g)
    labels[torch.randperm(M, generator=g)[: M // 9]] = -100       # some ignored rows (torch's ignore_index)
    labels = labels.cuda()
    n_valid = torch.tensor([int((labels != -100).sum())], dtype=torch.int32, device="cuda")
    scale = 1.0 / (M * 2)   # world = 2

    logits = torch.empty((M, N), dtype=dtype, device="cuda")
    d_ref = torch.full((M, N), float("nan"), dtype=dtype, device="cuda")
    loss_ref, db_ref = torch.zeros(1, device="cuda"), torch.zeros(N, device="cuda")
    ops.linear_fwd(x, w, b, logits, M=M, N=N, K=K, relu=False)
    ops.ce_loss_grad(logits, labels, d_ref, loss_ref, db_ref, torch.empty(ops.ce_scratch_floats(N), device="cuda"),
                     batch=M, num_classes=N, grad_scale=scale, n_valid=n_valid)

    d = torch.full((M, N), float("nan"), dtype=dtype, device="cuda")
    loss, db = torch.full((1,), float("nan"), device="cuda"), torch.full((N,), float("nan"), device="cuda")
    scratch = torch.empty(ops.linear_ce_scratch_floats(N), device="cuda")
    ops.linear_ce(x, w, b, labels, d, loss, db, scratch, M=M, N=N, K=K, grad_scale=scale, n_valid=n_valid)
    torch.cuda.synchronize()

    assert torch.isfinite(d.float()).all()
    assert torch.count_nonzero(d[labels == -100]).item() == 0
    torch.testing.assert_close(db.double(), d.double().sum(0), rtol=1e-4, atol=1e-7)   # sums of what IS stored
    if dtype == torch.float32:
        # same TF32 products, fp32 logits in both paths: only the order of the fp32 sums differs
        assert abs(loss.item() - loss_ref.item()) <= 1e-5 * max(1.0, abs(loss_ref.item()))
        torch.testing.assert_close(d, d_ref, rtol=1e-4, atol=1e-10)
    else:
        # bf16 operands are exact in the tensor core, accumulation is fp32: the fused kernel's logits are the fp64
        # ones to ~1e-6, so its gradient is torch's fp64 gradient rounded once to bf16 ...
        z = (x.double() @ w.double().t() + b.double()).requires_grad_(True)
        ref = F.cross_entropy(z, labels)
        ref.backward()
        assert abs(loss.item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item())), (loss.item(), ref.item())
        torch.testing.assert_close(d.double(), z.grad * (M * scale), rtol=8e-3, atol=1e-9)
        # ... while the two-kernel path first rounds every logit to bf16 (|dz| <= 2^-9 |z|): close, not equal
        assert abs(loss.item() - loss_ref.item()) <= 2e-3 * max(1.0, abs(loss_ref.item()))
        assert (d.float() - d_ref.float()).norm() <= 2e-2 * d_ref.float().norm()
    # deterministic, and `accumulate` adds onto loss / bias gradient
    loss2, db2 = loss.clone(), db.clone()
    ops.linear_ce(x, w, b, labels, d, loss2, db2, scratch, M=M, N=N, K=K, grad_scale=scale, n_valid=n_valid,
                  accumulate=True)
    torch.cuda.synchronize()
    assert torch.equal(loss2, loss + loss) and torch.equal(db2, db + db)


def test_linear_ce_rejects_more_than_1024_classes(ops):
    from b200step.lib import B200Error

    x = torch.zeros(128, 64, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(1032, 64, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(B200Error, match="1024 classes"):
        ops.linear_ce(x, w, torch.zeros(1032, device="cuda"), torch.zeros(128, dtype=torch.int64, device="cuda"),
                      torch.zeros(128, 1032, dtype=torch.bfloat16, device="cuda"), torch.zeros(1, device="cuda"), None,
                      torch.zeros(ops.linear_ce_scratch_floats(1024), device="cuda"), M=128, N=1032, K=64,
                      grad_scale=1.0)


# ------------------------------------------------------------ layer-0 dgrad + embedding backward in one kernel
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
@pytest.mark.parametrize("M,N,nc,rows,dim", [
    (4096, 512, 26, 1000, 16),      # T0
    (1000, 64, 26, 50, 16),         # small test model, ragged batch
    (2048, 1024, 26, 4096, 64),     # T1 row shape (dim 64 = two 32-column chunks per table row)
    (333, 128, 40, 17, 8),          # four tables per 32-column chunk, > 32 tables
    (512, 64, 4, 9, 4),
])
def test_linear_dgrad_emb_matches_dgrad_then_scatter(ops, dtype, M, N, nc, rows, dim):
    """b200_linear_dgrad_emb against b200_linear_dgrad + b200_emb_bwd and against torch index_add_ in fp64."""
    K = nc * dim
    kpad = (K + 13 + 63) // 64 * 64                       # the step's padded layer-0 width (dense + pad behind)
    dy = _rand((M, N), dtype, 51, 0.05)
    w = _rand((N, kpad), dtype, 52, 0.1)
    g = torch.Generator().manual_seed(53)
    idx = torch.randint(0, rows, (M, nc), generator=g, dtype=torch.int32).cuda()
    ref = torch.zeros(nc * rows * dim, device="cuda")
    dx = torch.empty(M, kpad, dtype=dtype, device="cuda")
    ops.linear_dgrad(dy, w, None, dx, M=M, N=N, K=kpad, relu=False)
    ops.emb_bwd(dx, idx, ref, batch=M, num_cat=nc, table_rows=rows, embed_dim=dim)
    got = torch.zeros_like(ref)
    ops.linear_dgrad_emb(dy, w, idx, got, M=M, N=N, num_cat=nc, table_rows=rows, embed_dim=dim)
    torch.cuda.synchronize()
    exact = torch.zeros(nc, rows, dim, dtype=torch.float64, device="cuda")
    dx64 = (dy.double() @ w.double())[:, :K].view(M, nc, dim)
    for t in range(nc):
        exact[t].index_add_(0, idx[:, t].long(), dx64[:, t])
    exact = exact.flatten()
    scale = exact.abs().max().item()
    # the fused kernel adds fp32 accumulators (bf16: exact products; f32: TF32 products) -> close to fp64;
    # the two-kernel path rounds dx to the storage dtype first
    tol = 2e-5 if dtype == torch.bfloat16 else 2e-3
    assert (got.double() - exact).abs().max().item() <= tol * scale
    assert (ref.double() - exact).abs().max().item() <= (1e-2 if dtype == torch.bfloat16 else 2e-3) * scale
    # accumulates onto what is there
    ops.linear_dgrad_emb(dy, w, idx, got, M=M, N=N, num_cat=nc, table_rows=rows, embed_dim=dim)
    torch.cuda.synchronize()
    assert (got.double() - 2 * exact).abs().max().item() <= 2 * tol * scale


def test_linear_dgrad_emb_rejects_odd_embedding_width(ops):
    from b200step.lib import B200Error

    with pytest.raises(B200Error, match="power-of-two"):
        ops.linear_dgrad_emb(torch.zeros(128, 64, dtype=torch.bfloat16, device="cuda"),
                             torch.zeros(64, 64, dtype=torch.bfloat16, device="cuda"),
                             torch.zeros(128, 4, dtype=torch.int32, device="cuda"),
                             torch.zeros(4 * 10 * 12, device="cuda"),
                             M=128, N=64, num_cat=4, table_rows=10, embed_dim=12)


# ----------------------------------------------------------------------------- 3xTF32: fp32-grade linear layers
def test_split_tf32_layouts_and_exactness(ops):
    R, C = 300, 96
    x = _rand((R, C + 8), torch.float32, 61)[:, :C]          # a strided view: ld > cols
    kcat = torch.full((R, 3 * C), float("nan"), device="cuda")
    rst = torch.full((3 * R, C), float("nan"), device="cuda")
    ops.split_tf32(x, kcat, ops.SPLIT_LEFT, rst, ops.SPLIT_RIGHT, rows=R, cols=C)
    torch.cuda.synchronize()
    hi, lo = kcat[:, :C], kcat[:, 2 * C:]
    assert torch.equal(kcat[:, C:2 * C], hi)                                  # (hi, hi, lo)
    assert torch.equal(hi + lo, x)                                            # the split is exact
    assert torch.count_nonzero(hi.view(torch.int32) & 0x1FFF).item() == 0     # hi is a tf32 number
    assert (lo.abs() <= hi.abs() * 2.0 ** -11 + 1e-45).all()
    assert torch.equal(rst[:R], hi) and torch.equal(rst[R:2 * R], lo) and torch.equal(rst[2 * R:], hi)   # (hi, lo, hi)


@pytest.mark.parametrize("M,N,K", [(4096, 512, 448), (333, 1000, 512)])
def test_three_pass_tf32_gemm_is_fp32_grade(ops, M, N, K):
    """[X_hi | X_hi | X_lo] . [W_hi | W_lo | W_hi]^T through the ordinary TF32 forward kernel: error ~1e-6 of the
    output scale, against ~1e-3 for the single pass."""
    x = _rand((M, K), torch.float32, 62)
    w = _rand((N, K), torch.float32, 63, 0.05)
    xk = torch.empty(M, 3 * K, device="cuda")
    wk = torch.empty(N, 3 * K, device="cuda")
    ops.split_tf32(x, xk, ops.SPLIT_LEFT, None, 0, rows=M, cols=K)
    ops.split_tf32(w, wk, ops.SPLIT_RIGHT, None, 0, rows=N, cols=K)
    y3 = torch.empty(M, N, device="cuda")
    y1 = torch.empty(M, N, device="cuda")
    ops.linear_fwd(xk, wk, None, y3, M=M, N=N, K=3 * K, relu=False)
    ops.linear_fwd(x, w, None, y1, M=M, N=N, K=K, relu=False)
    torch.cuda.synchronize()
    ref = x.double() @ w.double().t()
    scale = ref.abs().max().item()
    e3 = (y3.double() - ref).abs().max().item() / scale
    e1 = (y1.double() - ref).abs().max().item() / scale
    print(f"3xTF32 err {e3:.2e}, single TF32 err {e1:.2e}")
    assert e3 <= 2e-5 and e1 > 20 * e3     # (the tensor core's own fp32 accumulation is what is left)
