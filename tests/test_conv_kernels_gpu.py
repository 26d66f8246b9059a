"""Kernel-level parity of the convolutional path (csrc/conv.cu + the tcgen05 GEMMs it feeds) against torch's own
operators on the CPU: F.unfold / F.conv2d / F.batch_norm / F.max_pool2d with autograd, and torchvision's transform
restatement of oracle/resnet_step.py.  Data movement (augment in fp32, im2col, max-pool) is compared bit for bit;
floating-point reductions with the tolerance written at the assert.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from b200step import ops as _ops
    from b200step.lib import get_lib

    get_lib()
    return _ops


def _nhwc(t):      # NCHW cpu -> NHWC cuda matrix-friendly contiguous
    return t.permute(0, 2, 3, 1).contiguous().cuda()


def _nchw(t, B, H, W, C):
    return t.view(B, H, W, C).permute(0, 3, 1, 2).float().cpu()


GEOMS = [
    # B, H, C, k, stride, pad          (square images)
    (4, 32, 3, 7, 2, 3),               # conv1 of the ResNet stem
    (3, 8, 64, 3, 1, 1),               # layer1
    (3, 8, 64, 3, 2, 1),               # layer2.0.conv1
    (3, 8, 64, 1, 2, 0),               # layer2.0.downsample
    (5, 2, 256, 3, 2, 1),              # layer4.0.conv1: 2x2 -> 1x1
    (2, 1, 512, 3, 1, 1),              # layer4: a 3x3 filter over a single pixel
    (2, 7, 8, 3, 2, 1),                # odd size
]


def _col_reference(x, k, stride, pad):
    """F.unfold orders a patch (c, ky, kx); the kernels use (ky, kx, c)."""
    B, C, H, W = x.shape
    u = F.unfold(x, k, padding=pad, stride=stride)                 # [B, C*k*k, L]
    L = u.shape[-1]
    return u.view(B, C, k, k, L).permute(0, 4, 2, 3, 1).reshape(B * L, k * k * C)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad", GEOMS)
def test_im2col_is_unfold_bit_exact(ops, dtype, B, H, C, k, stride, pad):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, C, H, H, generator=g).to(dtype)
    ref = _col_reference(x.float(), k, stride, pad).to(dtype)
    Ho = ops.conv_out_size(H, k, stride, pad)
    K = k * k * C
    Kp = (K + 7) // 8 * 8
    col = torch.full((B * Ho * Ho, Kp), 7.0, dtype=dtype, device="cuda")
    ops.im2col(_nhwc(x), col, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad)
    torch.cuda.synchronize()
    assert torch.equal(col[:, :K].cpu(), ref)
    assert not col[:, K:].any()                                   # the GEMM's K padding is zero-filled


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad", [g for g in GEOMS if g[2] % 4 == 0])
def test_col2im_is_the_adjoint_of_im2col(ops, dtype, B, H, C, k, stride, pad):
    g = torch.Generator().manual_seed(2)
    Ho = ops.conv_out_size(H, k, stride, pad)
    K = k * k * C
    dcol = torch.randn(B * Ho * Ho, K, generator=g).to(dtype)
    add = torch.randn(B, C, H, H, generator=g).to(dtype)
    x = torch.zeros(B, C, H, H, requires_grad=True)
    (_col_reference(x, k, stride, pad) * dcol.float()).sum().backward()
    dx = torch.empty(B * H * H, C, dtype=dtype, device="cuda")
    for use_add in (False, True):
        ref = x.grad + (add.float() if use_add else 0)
        ops.col2im(dcol.cuda(), _nhwc(add) if use_add else None, dx, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride,
                   pad=pad)
        got = _nchw(dx, B, H, H, C)
        # <= 9 (49) terms summed in fp32 in another order than autograd's; bf16: one final rounding
        tol = 1e-5 if dtype == torch.float32 else 2e-2
        assert torch.allclose(got, ref, rtol=tol, atol=tol), (got - ref).abs().max()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["tf32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad,Cout", [
    (64, 32, 3, 7, 2, 3, 64), (32, 8, 64, 3, 1, 1, 64), (32, 8, 64, 3, 2, 1, 128), (32, 8, 64, 1, 2, 0, 128),
    (32, 2, 256, 3, 2, 1, 512), (32, 1, 512, 3, 1, 1, 512), (16, 4, 128, 3, 1, 1, 128)])
def test_convolution_forward_and_backward(ops, dtype, B, H, C, k, stride, pad, Cout):
    """im2col + tcgen05 GEMMs + col2im == F.conv2d and its autograd (the GEMM shapes N = 64..512, K = 152..4608)."""
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, C, H, H, generator=g).to(dtype)
    w = (torch.randn(Cout, C, k, k, generator=g) / (C * k * k) ** 0.5).to(dtype)
    xr = x.float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    yr = F.conv2d(xr, wr, stride=stride, padding=pad)
    dyr = torch.randn(yr.shape, generator=g).to(dtype)
    yr.backward(dyr.float())

    from b200step.convnet import filter_to_gemm, gemm_to_filter

    Ho = ops.conv_out_size(H, k, stride, pad)
    M, K = B * Ho * Ho, k * k * C
    Kp = (K + 7) // 8 * 8
    wg = filter_to_gemm(w, Kp).cuda()
    col = torch.empty(M, Kp, dtype=dtype, device="cuda")
    y = torch.empty(M, Cout, dtype=dtype, device="cuda")
    xg = _nhwc(x)
    ops.im2col(xg, col, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad)
    ops.linear_fwd(col, wg, None, y, M=M, N=Cout, K=Kp, relu=False)
    rt, at = (2e-2, 2e-2) if dtype == torch.bfloat16 else (5e-3, 5e-3)
    got = _nchw(y, B, Ho, Ho, Cout)
    assert torch.allclose(got, yr.detach(), rtol=rt, atol=at), (got - yr.detach()).abs().max()

    dy = _nhwc(dyr)
    dw = torch.empty(Cout, Kp, dtype=torch.float32, device="cuda")
    ops.linear_wgrad(dy.view(M, Cout), col, dw, M=M, N=Cout, K=Kp)
    gotw = gemm_to_filter(dw.cpu(), (Cout, C, k, k))
    scale = wr.grad.abs().max().item()
    assert (gotw - wr.grad).abs().max().item() <= rt * scale, ((gotw - wr.grad).abs().max().item(), scale)
    if C % 4 == 0:
        dcol = torch.empty(M, Kp, dtype=dtype, device="cuda")
        dx = torch.empty(B * H * H, C, dtype=dtype, device="cuda")
        ops.linear_dgrad(dy.view(M, Cout), wg, None, dcol, M=M, N=Cout, K=Kp, relu=False)
        ops.col2im(dcol, None, dx, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad)
        gotx = _nchw(dx, B, H, H, C)
        scale = xr.grad.abs().max().item()
        assert (gotx - xr.grad).abs().max().item() <= 2 * rt * scale


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("B,H,C,relu,with_res", [(32, 8, 64, True, False), (32, 4, 128, True, True),
                                                 (16, 2, 256, False, False), (48, 1, 512, True, True),
                                                 (1024, 16, 64, True, False), (3, 3, 64, True, True)])
def test_batchnorm_training_forward_backward(ops, dtype, B, H, C, relu, with_res):
    """nn.BatchNorm2d(train) (+ residual, + ReLU) and its backward against F.batch_norm under autograd, including the
    running-statistics update (momentum 0.1, unbiased variance) and num_batches_tracked."""
    g = torch.Generator().manual_seed(4)
    y = (torch.randn(B, C, H, H, generator=g) * 1.7 + 0.3).to(dtype)
    res = torch.randn(B, C, H, H, generator=g).to(dtype) if with_res else None
    gamma = torch.rand(C, generator=g) + 0.5
    beta = torch.randn(C, generator=g) * 0.1
    rm0, rv0 = torch.randn(C, generator=g) * 0.1, torch.rand(C, generator=g) + 0.5
    dout = torch.randn(B, C, H, H, generator=g).to(dtype)

    yr = y.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm, rv = rm0.clone(), rv0.clone()
    o = F.batch_norm(yr, rm, rv, gr, br, training=True, momentum=0.1, eps=1e-5)
    if res is not None:
        o = o + res.float()
    if relu:
        o = F.relu(o)
    o_store = o.detach().to(dtype)
    o.backward(dout.float())

    M = B * H * H
    f32 = dict(dtype=torch.float32, device="cuda")
    yg, out = _nhwc(y).view(M, C), torch.empty(M, C, dtype=dtype, device="cuda")
    mean, invstd = torch.empty(C, **f32), torch.empty(C, **f32)
    rmg, rvg, nb = rm0.cuda(), rv0.cuda(), torch.zeros(1, dtype=torch.int64, device="cuda")
    scratch = torch.zeros(ops.bn_scratch_floats(C), **f32)
    for _ in range(2):     # the second call proves the ticket / scratch are left ready
        rmg.copy_(rm0); rvg.copy_(rv0)
        ops.bn_fwd_stats(yg, mean, invstd, rmg, rvg, nb, scratch, M=M, C=C)
    resg = _nhwc(res).view(M, C) if res is not None else None
    ops.bn_fwd_apply(yg, resg, out, mean, invstd, gamma.cuda(), beta.cuda(), M=M, C=C, relu=relu)
    torch.cuda.synchronize()
    assert int(nb.item()) == 2
    mu_ref = y.float().mean(dim=(0, 2, 3))
    var_ref = y.float().var(dim=(0, 2, 3), unbiased=False)
    assert torch.allclose(mean.cpu(), mu_ref, rtol=1e-5, atol=1e-6)
    assert torch.allclose(invstd.cpu(), torch.rsqrt(var_ref + 1e-5), rtol=1e-5)
    assert torch.allclose(rmg.cpu(), rm, rtol=1e-5, atol=1e-6) and torch.allclose(rvg.cpu(), rv, rtol=1e-5, atol=1e-6)
    got = _nchw(out, B, H, H, C)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert torch.allclose(got, o_store.float(), rtol=tol, atol=tol), (got - o_store.float()).abs().max()

    dg, db = torch.empty(C, **f32), torch.empty(C, **f32)
    dy = torch.empty(M, C, dtype=dtype, device="cuda")
    dz = torch.empty(M, C, dtype=dtype, device="cuda")
    doutg = _nhwc(dout).view(M, C)
    mask_src = _nhwc(o_store).view(M, C) if relu else None       # the oracle's stored output decides the ReLU mask
    ops.bn_bwd_reduce(doutg, mask_src, yg, mean, invstd, dg, db, scratch, M=M, C=C)
    ops.bn_bwd_apply(doutg, mask_src, yg, dy, dz, mean, invstd, gamma.cuda(), dg, db, M=M, C=C)
    torch.cuda.synchronize()
    s = 1e-4 if dtype == torch.float32 else 2e-2
    assert torch.allclose(db.cpu(), br.grad, rtol=s, atol=s * br.grad.abs().max().item())
    assert torch.allclose(dg.cpu(), gr.grad, rtol=s, atol=s * gr.grad.abs().max().item())
    goty = _nchw(dy, B, H, H, C)
    assert (goty - yr.grad).abs().max().item() <= s * max(1.0, yr.grad.abs().max().item())
    dz_ref = dout.float() * (o_store.float() > 0) if relu else dout.float()
    assert torch.equal(_nchw(dz, B, H, H, C), dz_ref.to(dtype).float())


def test_batchnorm_statistics_are_run_to_run_identical(ops):
    y = torch.randn(65536, 64, device="cuda")
    f32 = dict(dtype=torch.float32, device="cuda")
    scratch = torch.zeros(ops.bn_scratch_floats(64), **f32)
    outs = []
    for _ in range(3):
        mean, invstd = torch.empty(64, **f32), torch.empty(64, **f32)
        ops.bn_fwd_stats(y, mean, invstd, None, None, None, scratch, M=65536, C=64)
        outs.append((mean.clone(), invstd.clone()))
    assert all(torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1]) for o in outs[1:])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("B,H,C", [(8, 16, 64), (3, 7, 64), (2, 2, 8)])
def test_maxpool_matches_torch_bit_exact(ops, dtype, B, H, C):
    g = torch.Generator().manual_seed(5)
    x = F.relu(torch.randn(B, C, H, H, generator=g)).to(dtype)      # post-ReLU: many exact ties at zero
    xr = x.float().requires_grad_(True)
    o = F.max_pool2d(xr, 3, 2, 1)
    Ho = o.shape[-1]
    dout = torch.randn(o.shape, generator=g).to(dtype)
    o.backward(dout.float())
    out = torch.empty(B * Ho * Ho, C, dtype=dtype, device="cuda")
    taps = torch.empty(B * Ho * Ho, C, dtype=torch.uint8, device="cuda")
    dx = torch.empty(B * H * H, C, dtype=dtype, device="cuda")
    ops.maxpool_fwd(_nhwc(x), out, taps, B=B, H=H, W=H, C=C)
    ops.maxpool_bwd(_nhwc(dout), taps, dx, B=B, H=H, W=H, C=C)
    assert torch.equal(_nchw(out, B, Ho, Ho, C), o.detach())
    ref = xr.grad
    got = _nchw(dx, B, H, H, C)
    if dtype == torch.float32:
        assert torch.allclose(got, ref, rtol=1e-6, atol=1e-6)       # up to 4 windows summed per pixel
    else:
        assert torch.allclose(got, ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_avgpool(ops, dtype):
    B, HW, C = 6, 4, 512
    g = torch.Generator().manual_seed(6)
    x = torch.randn(B, HW, C, generator=g).to(dtype)
    out = torch.empty(B, C, dtype=dtype, device="cuda")
    ops.avgpool_fwd(x.cuda(), out, B=B, HW=HW, C=C)
    assert torch.allclose(out.float().cpu(), x.float().mean(dim=1), rtol=1e-2 if dtype == torch.bfloat16 else 1e-6,
                          atol=1e-2 if dtype == torch.bfloat16 else 1e-6)
    d = torch.randn(B, C, generator=g).to(dtype)
    dx = torch.empty(B * HW, C, dtype=dtype, device="cuda")
    ops.avgpool_bwd(d.cuda(), dx, B=B, HW=HW, C=C)
    ref = (d.float() / HW).to(dtype)[:, None, :].expand(B, HW, C).reshape(B * HW, C)
    assert torch.equal(dx.cpu(), ref)


def test_augment_is_bit_identical_to_the_torchvision_pipeline(ops):
    """K-aug in fp32 against the reference's own transform object (task.py:250-255) run by torchvision on PIL images
    with the same random draws; the bf16 output is the rounding of the fp32 one."""
    from b200step.convnet import draw_augmentation, make_synthetic_images
    from oracle.resnet_step import augment_with_torchvision

    ds = make_synthetic_images(96, seed=3)
    torch.manual_seed(11)
    draws = draw_augmentation(96)
    ref = augment_with_torchvision(ds.images, seed=11)                   # [n, 3, 32, 32] fp32
    rows = torch.arange(96, dtype=torch.int64).flip(0).contiguous()       # any sampler order
    mean3 = torch.tensor([0.4914, 0.4822, 0.4465])
    std3 = torch.tensor([0.2023, 0.1994, 0.2010])
    x = torch.empty(96 * 32 * 32, 3, dtype=torch.float32, device="cuda")
    y = torch.empty(96, dtype=torch.int64, device="cuda")
    ops.image_augment(ds.images.cuda(), ds.labels.cuda(), rows.cuda(), None, draws.flip(0).contiguous().cuda(), mean3,
                      std3, x, y, batch=96, H=32, W=32)
    got = x.view(96, 32, 32, 3).permute(0, 3, 1, 2).cpu()
    assert torch.equal(got, ref.flip(0))
    assert torch.equal(y.cpu(), ds.labels.flip(0))
    xb = torch.empty(96 * 32 * 32, 3, dtype=torch.bfloat16, device="cuda")
    ops.image_augment(ds.images.cuda(), ds.labels.cuda(), rows.cuda(), None, draws.flip(0).contiguous().cuda(), mean3,
                      std3, xb, None, batch=96, H=32, W=32)
    assert torch.equal(xb.cpu(), x.cpu().to(torch.bfloat16))


def test_augment_reports_a_bad_sample_index(ops):
    from b200step.convnet import make_synthetic_images

    ds = make_synthetic_images(8, seed=0)
    word = ops.status_word()
    word.zero_()
    rows = torch.tensor([0, 1, 99, 3], dtype=torch.int64, device="cuda")
    x = torch.empty(4 * 32 * 32, 3, dtype=torch.float32, device="cuda")
    ops.image_augment(ds.images.cuda(), ds.labels.cuda(), rows, None, None, torch.zeros(3), torch.ones(3), x, None,
                      batch=4, H=32, W=32)
    torch.cuda.synchronize()
    assert int(word.item()) & ops.STATUS_BAD_ROW_INDEX
    word.zero_()


IMPLICIT_GEOMS = [
    # B, H, C, k, stride, pad, Cout
    (32, 8, 64, 3, 1, 1, 64), (32, 8, 64, 3, 2, 1, 128), (32, 8, 64, 1, 2, 0, 128), (32, 2, 256, 3, 2, 1, 512),
    (32, 1, 512, 3, 1, 1, 512), (16, 4, 128, 3, 1, 1, 128), (3, 5, 64, 3, 1, 1, 72), (1024, 8, 64, 3, 1, 1, 64),
    (2, 3, 128, 3, 2, 1, 64),
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["tf32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad,Cout", IMPLICIT_GEOMS)
def test_implicit_gemm_convolution_forward(ops, dtype, B, H, C, k, stride, pad, Cout):
    """b200_conv_fwd (A tiles by im2col-mode TMA, no col matrix) against the explicit im2col + GEMM path, which must
    agree BIT FOR BIT (same operand bytes reach the same MMAs in the same order), and against F.conv2d."""
    from b200step.convnet import filter_to_gemm

    g = torch.Generator().manual_seed(8)
    x = torch.randn(B, C, H, H, generator=g).to(dtype)
    w = (torch.randn(Cout, C, k, k, generator=g) / (C * k * k) ** 0.5).to(dtype)
    ref = F.conv2d(x.float(), w.float(), stride=stride, padding=pad)
    Ho = ops.conv_out_size(H, k, stride, pad)
    M, K = B * Ho * Ho, k * k * C
    wg = filter_to_gemm(w, K).cuda()
    xg = _nhwc(x)
    col = torch.empty(M, K, dtype=dtype, device="cuda")
    y_explicit = torch.empty(M, Cout, dtype=dtype, device="cuda")
    ops.im2col(xg, col, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad)
    ops.linear_fwd(col, wg, None, y_explicit, M=M, N=Cout, K=K, relu=False)
    y = torch.full((M, Cout), 9.0, dtype=dtype, device="cuda")
    # deterministic: fp32 outputs may otherwise split the contraction of a small convolution (atomics, see the split test)
    ops.conv_fwd(xg, wg, y, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad, Cout=Cout, deterministic=True)
    torch.cuda.synchronize()
    assert torch.equal(y, y_explicit)
    rt = 2e-2 if dtype == torch.bfloat16 else 5e-3
    assert torch.allclose(_nchw(y, B, Ho, Ho, Cout), ref, rtol=rt, atol=rt)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["tf32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad,Cout", IMPLICIT_GEOMS)
def test_implicit_gemm_filter_gradient(ops, dtype, B, H, C, k, stride, pad, Cout):
    """b200_conv_wgrad (B tiles by im2col-mode TMA) against the explicit dy^T . col GEMM — bit for bit in deterministic
    mode — and against autograd."""
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, C, H, H, generator=g).to(dtype)
    wr = (torch.randn(Cout, C, k, k, generator=g) / (C * k * k) ** 0.5).requires_grad_(True)
    yr = F.conv2d(x.float(), wr, stride=stride, padding=pad)
    dyr = torch.randn(yr.shape, generator=g).to(dtype)
    yr.backward(dyr.float())
    from b200step.convnet import gemm_to_filter

    Ho = ops.conv_out_size(H, k, stride, pad)
    M, K = B * Ho * Ho, k * k * C
    xg, dy = _nhwc(x), _nhwc(dyr).view(M, Cout)
    col = torch.empty(M, K, dtype=dtype, device="cuda")
    ops.im2col(xg, col, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad)
    dw_explicit = torch.empty(Cout, K, dtype=torch.float32, device="cuda")
    ops.linear_wgrad(dy, col, dw_explicit, M=M, N=Cout, K=K, deterministic=True)
    dw = torch.full((Cout, K), 5.0, dtype=torch.float32, device="cuda")
    ops.conv_wgrad(dy, xg, dw, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad, Cout=Cout, deterministic=True)
    torch.cuda.synchronize()
    assert torch.equal(dw, dw_explicit)
    dw2 = torch.full((Cout, K), 5.0, dtype=torch.float32, device="cuda")      # default: batch dimension split, atomics
    ops.conv_wgrad(dy, xg, dw2, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad, Cout=Cout)
    rt = 2e-2 if dtype == torch.bfloat16 else 5e-3
    scale = wr.grad.abs().max().item()
    for t in (dw, dw2):
        got = gemm_to_filter(t.cpu(), (Cout, C, k, k))
        assert (got - wr.grad).abs().max().item() <= rt * scale


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["tf32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,pad,Cout", [(32, 8, 64, 3, 1, 64), (16, 4, 128, 3, 1, 128), (32, 1, 512, 3, 1, 512),
                                              (8, 2, 256, 3, 1, 256), (3, 5, 64, 3, 1, 128), (1024, 8, 64, 3, 1, 64),
                                              (4, 6, 72, 1, 0, 64)])
def test_implicit_gemm_input_gradient(ops, dtype, B, H, C, k, pad, Cout):
    """b200_conv_dgrad (stride 1: dy convolved with the flipped filter, A tiles by im2col-mode TMA over dy, B tiles from
    the forward's own filter matrix) against dgrad GEMM + col2im and against autograd."""
    from b200step.convnet import filter_to_gemm

    g = torch.Generator().manual_seed(10)
    w = (torch.randn(Cout, C, k, k, generator=g) / (C * k * k) ** 0.5).to(dtype)
    xr = torch.zeros(B, C, H, H, requires_grad=True)
    yr = F.conv2d(xr, w.float(), stride=1, padding=pad)
    dyr = torch.randn(yr.shape, generator=g).to(dtype)
    yr.backward(dyr.float())
    Ho = ops.conv_out_size(H, k, 1, pad)
    M, K = B * Ho * Ho, k * k * C
    wg = filter_to_gemm(w, K).cuda()
    dy = _nhwc(dyr).view(M, Cout)
    dcol = torch.empty(M, K, dtype=dtype, device="cuda")
    dx_explicit = torch.empty(B * H * H, C, dtype=dtype, device="cuda")
    ops.linear_dgrad(dy, wg, None, dcol, M=M, N=Cout, K=K, relu=False)
    ops.col2im(dcol, None, dx_explicit, B=B, H=H, W=H, C=C, kh=k, kw=k, stride=1, pad=pad)
    dx = torch.full((B * H * H, C), 3.0, dtype=dtype, device="cuda")
    ops.conv_dgrad(dy, wg, dx, B=B, H=H, W=H, C=C, kh=k, kw=k, pad=pad, Cout=Cout)
    torch.cuda.synchronize()
    rt = 2e-2 if dtype == torch.bfloat16 else 5e-3
    scale = xr.grad.abs().max().item()
    got = _nchw(dx, B, H, H, C)
    assert (got - xr.grad).abs().max().item() <= 2 * rt * scale
    # same products, fp32 accumulation in another order (one long contraction instead of 9 rounded partial sums)
    assert (dx.float() - dx_explicit.float()).abs().max().item() <= 2 * rt * scale


def test_batchnorm_backward_sums_two_incoming_gradients(ops):
    g = torch.Generator().manual_seed(12)
    M, C = 2048, 64
    f32 = dict(dtype=torch.float32, device="cuda")
    y, out = torch.randn(M, C, generator=g).cuda(), torch.randn(M, C, generator=g).cuda()
    a, b = torch.randn(M, C, generator=g).cuda(), torch.randn(M, C, generator=g).cuda()
    mean, invstd, gamma = torch.randn(C, generator=g).cuda(), (torch.rand(C, generator=g) + 0.5).cuda(), torch.ones(C, **f32)
    scr = torch.zeros(ops.bn_scratch_floats(C), **f32)
    res = []
    for d1, d2 in ((a + b, None), (a, b)):
        dg, db = torch.empty(C, **f32), torch.empty(C, **f32)
        dy, dz = torch.empty(M, C, **f32), torch.empty(M, C, **f32)
        ops.bn_bwd_reduce(d1, out, y, mean, invstd, dg, db, scr, M=M, C=C, dout2=d2)
        ops.bn_bwd_apply(d1, out, y, dy, dz, mean, invstd, gamma, dg, db, M=M, C=C, dout2=d2)
        res.append((dg, db, dy, dz))
    for u, v in zip(*res):
        assert torch.equal(u, v)
    # max-pool backward too
    B, H = 4, 8
    x = torch.relu(torch.randn(B * H * H, C, generator=g)).cuda()
    o = torch.empty(B * 16, C, **f32)
    taps = torch.empty(B * 16, C, dtype=torch.uint8, device="cuda")
    ops.maxpool_fwd(x, o, taps, B=B, H=H, W=H, C=C)
    d1, d2 = torch.randn(B * 16, C, generator=g).cuda(), torch.randn(B * 16, C, generator=g).cuda()
    r1, r2 = torch.empty(B * H * H, C, **f32), torch.empty(B * H * H, C, **f32)
    ops.maxpool_bwd(d1 + d2, taps, r1, B=B, H=H, W=H, C=C)
    ops.maxpool_bwd(d1, taps, r2, B=B, H=H, W=H, C=C, dout2=d2)
    assert torch.equal(r1, r2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["tf32", "bf16"])
@pytest.mark.parametrize("B,H,C,k,stride,pad,Cout", [(1024, 1, 512, 3, 1, 1, 512), (1024, 2, 256, 3, 1, 1, 256),
                                                     (1024, 2, 256, 3, 2, 1, 512), (37, 2, 128, 3, 1, 1, 192)])
def test_split_contraction_of_small_convolutions(ops, dtype, B, H, C, k, stride, pad, Cout):
    """With a workspace, b200_conv_fwd / b200_conv_dgrad split the contraction of a small-M convolution over the idle
    SMs (partials combined by fp32 atomics).  Same products: the result equals the unsplit one up to the order of the
    fp32 additions (and one final rounding for bf16)."""
    from b200step.convnet import filter_to_gemm

    g = torch.Generator().manual_seed(13)
    x = torch.randn(B, C, H, H, generator=g).to(dtype)
    w = (torch.randn(Cout, C, k, k, generator=g) / (C * k * k) ** 0.5).to(dtype)
    Ho = ops.conv_out_size(H, k, stride, pad)
    M, K = B * Ho * Ho, k * k * C
    wg, xg = filter_to_gemm(w, K).cuda(), _nhwc(x)
    ws = torch.full((37 * 256 * 256,), 7.0, dtype=torch.float32, device="cuda")
    y0 = torch.empty(M, Cout, dtype=dtype, device="cuda")
    y1 = torch.full((M, Cout), 9.0, dtype=dtype, device="cuda")
    kw = dict(B=B, H=H, W=H, C=C, kh=k, kw=k, stride=stride, pad=pad, Cout=Cout)
    ops.conv_fwd(xg, wg, y0, deterministic=True, **kw)
    for _ in range(2):                                   # twice: the workspace is re-zeroed by the call itself
        ops.conv_fwd(xg, wg, y1, ws=ws, **kw)
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
    assert torch.allclose(y1.float(), y0.float(), rtol=tol, atol=tol), (y1.float() - y0.float()).abs().max()
    if stride == 1:
        dy = torch.randn(M, Cout, generator=g).to(dtype).cuda()
        d0 = torch.empty(B * H * H, C, dtype=dtype, device="cuda")
        d1 = torch.full((B * H * H, C), 9.0, dtype=dtype, device="cuda")
        kd = dict(B=B, H=H, W=H, C=C, kh=k, kw=k, pad=pad, Cout=Cout)
        ops.conv_dgrad(dy, wg, d0, deterministic=True, **kd)
        ops.conv_dgrad(dy, wg, d1, ws=ws, **kd)
        s = d0.float().abs().max().item()
        assert (d1.float() - d0.float()).abs().max().item() <= tol * max(1.0, s)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("M,C,relu,with_res,two", [(2048, 64, True, False, False), (512, 128, True, True, True),
                                                   (64, 256, False, False, False), (48, 512, True, True, True),
                                                   (262144, 64, True, False, False), (65536, 64, True, True, True),
                                                   (27, 64, True, True, False)])
def test_fused_batchnorm_launches_equal_the_two_launch_pairs(ops, dtype, M, C, relu, with_res, two):
    """b200_bn_fwd / b200_bn_bwd (statistics + apply in ONE launch, blocks waiting in-kernel for the last block's result)
    must reproduce the stats/apply and reduce/apply pairs bit for bit, repeatedly (the epoch / ticket words re-arm)."""
    g = torch.Generator().manual_seed(21)
    f32 = dict(dtype=torch.float32, device="cuda")
    y = (torch.randn(M, C, generator=g) * 1.3 + 0.2).to(dtype).cuda()
    res = torch.randn(M, C, generator=g).to(dtype).cuda() if with_res else None
    gamma, beta = (torch.rand(C, generator=g) + 0.5).cuda(), (torch.randn(C, generator=g) * 0.1).cuda()
    rm0, rv0 = (torch.randn(C, generator=g) * 0.1).cuda(), (torch.rand(C, generator=g) + 0.5).cuda()
    scr = torch.zeros(ops.bn_scratch_floats(C), **f32)
    mean, invstd = torch.empty(C, **f32), torch.empty(C, **f32)
    rm, rv, nb = rm0.clone(), rv0.clone(), torch.zeros(1, dtype=torch.int64, device="cuda")
    out = torch.empty(M, C, dtype=dtype, device="cuda")
    ops.bn_fwd_stats(y, mean, invstd, rm, rv, nb, scr, M=M, C=C)
    ops.bn_fwd_apply(y, res, out, mean, invstd, gamma, beta, M=M, C=C, relu=relu)
    for rep in range(3):
        mean2, invstd2 = torch.empty(C, **f32), torch.empty(C, **f32)
        rm2, rv2, nb2 = rm0.clone(), rv0.clone(), torch.zeros(1, dtype=torch.int64, device="cuda")
        out2 = torch.full((M, C), 3.0, dtype=dtype, device="cuda")
        ops.bn_fwd(y, res, out2, mean2, invstd2, gamma, beta, rm2, rv2, nb2, scr, M=M, C=C, relu=relu)
        torch.cuda.synchronize()
        assert torch.equal(mean, mean2) and torch.equal(invstd, invstd2) and torch.equal(out, out2), rep
        assert torch.equal(rm, rm2) and torch.equal(rv, rv2) and int(nb2.item()) == 1
    dout = torch.randn(M, C, generator=g).to(dtype).cuda()
    dout2 = torch.randn(M, C, generator=g).to(dtype).cuda() if two else None
    mask = out if relu else None
    dg, db = torch.empty(C, **f32), torch.empty(C, **f32)
    dy, dz = torch.empty(M, C, dtype=dtype, device="cuda"), torch.empty(M, C, dtype=dtype, device="cuda")
    ops.bn_bwd_reduce(dout, mask, y, mean, invstd, dg, db, scr, M=M, C=C, dout2=dout2)
    ops.bn_bwd_apply(dout, mask, y, dy, dz, mean, invstd, gamma, dg, db, M=M, C=C, dout2=dout2)
    for rep in range(3):
        dg2, db2 = torch.empty(C, **f32), torch.empty(C, **f32)
        dy2 = torch.full((M, C), 3.0, dtype=dtype, device="cuda")
        dz2 = torch.full((M, C), 3.0, dtype=dtype, device="cuda")
        ops.bn_bwd(dout, mask, y, dy2, dz2, mean, invstd, gamma, dg2, db2, scr, M=M, C=C, dout2=dout2)
        torch.cuda.synchronize()
        assert torch.equal(dg, dg2) and torch.equal(db, db2) and torch.equal(dz, dz2), rep
        # dy: the same formula in two kernels; ptxas contracts its multiply-adds differently (last-bit differences)
        tol = 1e-5 if dtype == torch.float32 else 1e-2
        assert torch.allclose(dy.float(), dy2.float(), rtol=tol, atol=tol * 0.1), rep
