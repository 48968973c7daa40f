"""GPU: each C-ABI kernel against a plain torch reference (fp64 where it is a contraction), forward and backward.
Tolerances are written beside each check."""
import math

import pytest
import torch
import torch.nn.functional as F

from oracle import bivae_oracle as O
from tests.helpers import rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import mmvae_b200
    return mmvae_b200.ops


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("M,N,K", [(128, 128, 16), (300, 70, 37), (7680, 512, 20), (512, 20, 7680), (97, 784, 512),
                                   (1000, 48, 32), (5, 3, 9), (128, 2048, 20), (7680, 20, 2048), (20, 2048, 7680)])
@pytest.mark.parametrize("a_k,b_k", [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_gemm_layouts(ops, M, N, K, a_k, b_k):
    A = _rand(M, K, seed=1)
    B = _rand(N, K, seed=2)
    bias = _rand(N, seed=3)
    a_store = A.contiguous() if a_k else A.t().contiguous()
    b_store = B.contiguous() if b_k else B.t().contiguous()
    ref = A.double() @ B.double().t()
    tol = 2e-6                                         # fp32 FMA / 3xTF32 accumulation vs fp64
    out = ops.gemm(a_store, b_store, None, M, N, K, a_k, b_k)
    assert rel_err(out, ref) < tol
    out = ops.gemm(a_store, b_store, bias, M, N, K, a_k, b_k, "relu")
    assert rel_err(out, torch.relu(ref + bias.double())) < tol
    out = ops.gemm(a_store, b_store, bias, M, N, K, a_k, b_k, "sigmoid")
    assert rel_err(out, torch.sigmoid(ref + bias.double())) < tol


@pytest.mark.parametrize("M,N,K", [(128, 512, 784), (128, 20, 2048), (20, 512, 128), (77, 53, 131), (1, 1, 1), (64, 64, 16),
                                   (65, 129, 33), (2048, 128, 1024), (128, 128, 128), (40, 128, 128), (3, 1000, 515)])
@pytest.mark.parametrize("a_k,b_k", [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_gemm_small_cluster_split_k(ops, M, N, K, a_k, b_k):
    """The one-launch cluster split-K kernel (gemm_small_kernel: K split over the CTAs of a thread-block cluster, partials
    reduced through distributed shared memory): every split factor 1..16, ragged M / N / K, unaligned leading dimensions, all
    four operand layouts, bias + activation; exact fp32 FMA against fp64; bit-identical on a second call (fixed reduction
    order); and the same shapes with the kernel switched off (limit 0) agree to rounding."""
    from mmvae_b200 import _lib
    lib = _lib.load()
    A, B, bias = _rand(M, K, seed=1, scale=K ** -0.5), _rand(N, K, seed=2), _rand(N, seed=3)
    a_s = A.contiguous() if a_k else A.t().contiguous()
    b_s = B.contiguous() if b_k else B.t().contiguous()
    ref = A.double() @ B.double().t()
    for act, fn in ((None, None), ("relu", torch.relu), ("sigmoid", torch.sigmoid)):
        n0 = lib.mmvae_gemm_tc_launch_count()
        out = ops.gemm(a_s, b_s, bias if act else None, M, N, K, a_k, b_k, act)
        want = fn(ref + bias.double()) if act else ref
        assert (out.double() - want).abs().max().item() <= 2e-6 * max(1.0, want.abs().max().item())
        again = ops.gemm(a_s, b_s, bias if act else None, M, N, K, a_k, b_k, act)
        assert torch.equal(out, again)
        assert lib.mmvae_gemm_tc_launch_count() == n0                     # neither call went to the tensor-core kernel
        old = lib.mmvae_gemm_small_set_limit(0)
        try:
            other = ops.gemm(a_s, b_s, bias if act else None, M, N, K, a_k, b_k, act)
        finally:
            lib.mmvae_gemm_small_set_limit(old)
        assert (out - other).abs().max().item() <= 4e-6 * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("M,N,K", [(7680, 784, 512), (7680, 512, 784), (512, 784, 7680), (7680, 512, 512)])
@pytest.mark.parametrize("a_k,b_k", [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_gemm_tensor_core_path_with_epilogue(ops, M, N, K, a_k, b_k):
    """The MLP-decoder shapes of C2 (above the 0.4 GMAC gate of gemm_tc.cu) in all four operand layouts, with the bias + ReLU /
    Sigmoid epilogue and without: these must run on gemm_tc_kernel (BN = 128 tiles, split-K for the K = 7680 one), which the
    launch counter asserts, and match fp64."""
    from mmvae_b200 import _lib
    lib = _lib.load()
    A, B = _rand(M, K, seed=1, scale=K ** -0.5), _rand(N, K, seed=2)      # unit-scale pre-activations: no saturated sigmoids
    bias = _rand(N, seed=3)
    a_s = A if a_k else A.t().contiguous()
    b_s = B if b_k else B.t().contiguous()
    ref = A.double() @ B.double().t()
    for act, fn in ((None, None), ("relu", torch.relu), ("sigmoid", torch.sigmoid)):
        n0 = lib.mmvae_gemm_tc_launch_count()
        out = ops.gemm(a_s, b_s, bias if act else None, M, N, K, a_k, b_k, act)
        assert lib.mmvae_gemm_tc_launch_count() == n0 + 1, "shape did not run on the tcgen05 kernel"
        r = fn(ref + bias.double()) if act else ref
        assert rel_err(out, r) < 2e-6, (act, rel_err(out, r))
    if a_k and b_k:                                     # the opt-in bf16x3 mode of the same kernel (debug bit 65536): 3 * 2^-18 per product
        old = lib.mmvae_gemm_tc_set_mode(65536 << 8)
        try:
            out = ops.gemm(a_s, b_s, bias, M, N, K, a_k, b_k, "relu")
        finally:
            lib.mmvae_gemm_tc_set_mode(old)
        assert rel_err(out, torch.relu(ref + bias.double())) < 1e-5


def test_colsum_and_act_bwd(ops):
    A = _rand(1000, 77, seed=4)
    assert rel_err(ops.colsum(A), A.double().sum(0)) < 1e-6
    y = torch.sigmoid(_rand(333, 5, seed=5))
    g = _rand(333, 5, seed=6)
    assert rel_err(ops.act_bwd(g, y, "sigmoid"), g * y * (1 - y)) < 1e-6
    y = torch.relu(_rand(333, 5, seed=7))
    assert torch.equal(ops.act_bwd(g, y, "relu"), g * (y > 0))


def test_linear_autograd(ops):
    x = _rand(50, 20, seed=1).requires_grad_()
    w = _rand(64, 20, seed=2, scale=0.2).requires_grad_()
    b = _rand(64, seed=3).requires_grad_()
    for act, fn in (("relu", torch.relu), ("sigmoid", torch.sigmoid), (None, lambda t: t)):
        y = ops.LinearAct.apply(x, w, b, act)
        gy = _rand(50, 64, seed=9)
        gx, gw, gb = torch.autograd.grad(y, (x, w, b), gy)
        xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
        yr = fn(xd @ wd.t() + bd)
        rx, rw, rb = torch.autograd.grad(yr, (xd, wd, bd), gy.double())
        assert rel_err(y, yr) < 2e-6
        assert rel_err(gx, rx) < 5e-6 and rel_err(gw, rw) < 5e-6 and rel_err(gb, rb) < 5e-6


@pytest.mark.parametrize("cin,cout,H,s,p,nchw_in", [(3, 32, 32, 2, 1, True), (32, 64, 16, 2, 1, False),
                                                    (128, 20, 4, 2, 0, False), (1, 8, 28, 2, 1, True)])
def test_conv2d_autograd(ops, cin, cout, H, s, p, nchw_in):
    N = 5
    x = _rand(N, cin, H, H, seed=1)
    if not nchw_in:
        x = x.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)        # NHWC memory, logical NCHW
    x.requires_grad_()
    w = _rand(cout, cin, 4, 4, seed=2, scale=0.1).requires_grad_()
    b = _rand(cout, seed=3).requires_grad_()
    y = ops.Conv2dAct.apply(x, w, b, s, p, "relu")
    xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
    yr = torch.relu(F.conv2d(xd, wd, bd, stride=s, padding=p))
    assert y.shape == yr.shape
    assert rel_err(y, yr) < 2e-6
    gy = _rand(*yr.shape, seed=4)
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), gy)
    rx, rw, rb = torch.autograd.grad(yr, (xd, wd, bd), gy.double())
    assert rel_err(gx, rx) < 5e-6 and rel_err(gw, rw) < 5e-6 and rel_err(gb, rb) < 5e-6


@pytest.mark.parametrize("cin,cout,H,s,p,act,nchw_out,N", [(20, 128, 1, 1, 0, "relu", False, None), (128, 64, 4, 2, 1, "relu", False, None),
                                                         (32, 3, 16, 2, 1, "sigmoid", True, None), (64, 32, 8, 2, 1, "relu", False, None),
                                                         (32, 1, 14, 2, 1, "sigmoid", True, None), (32, 4, 5, 2, 1, "relu", True, None),
                                                         (32, 3, 16, 2, 1, None, True, None), (16, 3, 8, 2, 1, "sigmoid", True, None),
                                                         (32, 1, 16, 2, 1, "sigmoid", True, None), (32, 3, 16, 2, 1, "relu", True, 1),
                                                         (128, 64, 4, 2, 1, "relu", False, 77), (64, 32, 8, 2, 1, "relu", False, 77),
                                                         (32, 32, 16, 2, 1, None, False, 5), (32, 64, 2, 2, 1, "relu", False, 131),
                                                         (128, 128, 8, 2, 1, "sigmoid", False, 3)])
def test_convT2d_autograd(ops, cin, cout, H, s, p, act, nchw_out, N):
    if N is None:
        N = 6 if cout > 4 or H != 16 else 301      # the direct path's persistent wgrad loops over images per CTA
    x = _rand(N, cin, H, H, seed=1).permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2).requires_grad_()
    w = _rand(cin, cout, 4, 4, seed=2, scale=0.1).requires_grad_()
    b = _rand(cout, seed=3).requires_grad_()
    y = ops.ConvT2dAct.apply(x, w, b, s, p, act, nchw_out, False)
    xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
    fn = {"relu": torch.relu, "sigmoid": torch.sigmoid, None: (lambda t: t)}[act]
    yr = fn(F.conv_transpose2d(xd, wd, bd, stride=s, padding=p))
    assert y.shape == yr.shape
    assert y.is_contiguous() == nchw_out or cout == 1
    # shapes on the second-generation kernels (convt_ts.cu) compute hi*hi + hi*mid + mid*hi on bf16 pieces: each product is
    # within 3 * 2^-18 = 1.1e-5 of exact (measured 3-5e-6 on the tensor); the 3xTF32 / SIMT paths are at 1e-6
    from mmvae_b200 import _lib
    ts = s == 2 and p == 1 and not nchw_out and _lib.load().mmvae_convt_ts_supported(H, H, cin, cout) == 1
    assert rel_err(y, yr) < (1e-5 if ts else 2e-6)
    gy = _rand(*yr.shape, seed=4)
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), gy)
    if act == "relu":
        # branch-consistent reference: the ReLU mask is the kernel's own (a unit whose pre-activation is within the forward
        # error of zero may land on either side of the kink; each such flip moves the gradient by an O(1) term, 1e-3 of the
        # norm at N = 77, which says nothing about the arithmetic being compared)
        pre = F.conv_transpose2d(xd, wd, bd, stride=s, padding=p)
        rx, rw, rb = torch.autograd.grad(pre, (xd, wd, bd), gy.double() * (y.detach() > 0))
    else:
        rx, rw, rb = torch.autograd.grad(yr, (xd, wd, bd), gy.double())
    tol = 1.5e-5 if ts else 5e-6
    l4w = (cin, cout, H, s, p) == (32, 3, 16, 2, 1) and nchw_out        # the image layer's weight gradient: bf16x3 (convt_l4w_ts_kernel)
    assert rel_err(gx, rx) < tol, "gx"
    assert rel_err(gw, rw) < (1.5e-5 if l4w else tol), "gw"
    assert rel_err(gb, rb) < 5e-6, "gb"


def test_convT_last_layer_tensor_core_forward_is_deterministic_and_matches_simt(ops):
    """convt_l4_tc.cu (tcgen05, col block kept on the SM) against the direct SIMT kernel it replaces and against fp64."""
    from mmvae_b200 import _lib
    lib = _lib.load()
    N = 333
    x = _rand(N, 16, 16, 32, seed=5).permute(0, 3, 1, 2)                # NHWC memory, logical NCHW
    w = _rand(32, 3, 4, 4, seed=6, scale=0.1)
    b = _rand(3, seed=7)
    y1 = ops.ConvT2dAct.apply(x, w, b, 2, 1, "sigmoid", True, False)
    y2 = ops.ConvT2dAct.apply(x, w, b, 2, 1, "sigmoid", True, False)
    assert torch.equal(y1, y2)                                          # fixed summation order
    old = lib.mmvae_gemm_tc_set_mode(512 << 8)                          # debug bit 512: force the SIMT kernel
    try:
        y_simt = ops.ConvT2dAct.apply(x, w, b, 2, 1, "sigmoid", True, False)
    finally:
        lib.mmvae_gemm_tc_set_mode(old)
    ref = torch.sigmoid(F.conv_transpose2d(x.double(), w.double(), b.double(), stride=2, padding=1))
    assert rel_err(y_simt, ref) < 2e-6 and rel_err(y1, ref) < 2e-6
    assert (y1 - y_simt).abs().max().item() < 2e-6


@pytest.mark.parametrize("N", [1, 7, 333, 1200])
def test_convT_last_layer_tensor_core_wgrad(ops, N):
    """The image layer's weight gradient: convt_l4w_ts_kernel (default: g transposed into tensor memory, bf16x3, chains of 12
    pixel tiles drained into registers), the fp32-FMA kernel (debug bit 32) and the first-generation tcgen05 kernel (bit 16384)
    against fp64 autograd and each other; ragged image counts per CTA at N = 333, several chains per CTA at N = 1200."""
    from mmvae_b200 import _lib
    lib = _lib.load()
    x = _rand(N, 16, 16, 32, seed=5).relu().permute(0, 3, 1, 2).requires_grad_()
    w = _rand(32, 3, 4, 4, seed=6, scale=0.1).requires_grad_()
    b = _rand(3, seed=7).requires_grad_()
    gy = _rand(N, 3, 32, 32, seed=8)
    xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
    yr = torch.sigmoid(F.conv_transpose2d(xd, wd, bd, stride=2, padding=1))
    rw, rb = torch.autograd.grad(yr, (wd, bd), gy.double())
    res = {}
    tols = {0: 1.5e-5, 32: 5e-6, 16384: 5e-6}          # 0: tensor-memory bf16x3 kernel (convt_ts.cu), 32: fp32 FMA, 16384: first-generation tcgen05
    for bit in (0, 32, 16384):
        old = lib.mmvae_gemm_tc_set_mode(bit << 8)
        try:
            y = ops.ConvT2dAct.apply(x, w, b, 2, 1, "sigmoid", True, False)
            res[bit] = torch.autograd.grad(y, (w, b), gy)
            again = torch.autograd.grad(ops.ConvT2dAct.apply(x, w, b, 2, 1, "sigmoid", True, False), (w, b), gy)
            assert torch.equal(res[bit][0], again[0]) and torch.equal(res[bit][1], again[1])     # deterministic
        finally:
            lib.mmvae_gemm_tc_set_mode(old)
        assert rel_err(res[bit][0], rw) < tols[bit] and rel_err(res[bit][1], rb) < 5e-6, bit
    assert rel_err(res[16384][0], res[32][0]) < 5e-6 and rel_err(res[0][0], res[32][0]) < 1.5e-5


def test_convT2d_chain_fused_relu_backward(ops):
    """Decoder_VAE_SVHN's chaining: each layer hands the layer below the gradient w.r.t. its PRE-activation."""
    N = 9
    z = _rand(N, 20, 1, 1, seed=1).requires_grad_()
    shapes = [(20, 128), (128, 64), (64, 32), (32, 3)]
    ws = [_rand(a, b, 4, 4, seed=10 + i, scale=0.1).requires_grad_() for i, (a, b) in enumerate(shapes)]
    bs = [_rand(b, seed=20 + i).requires_grad_() for i, (a, b) in enumerate(shapes)]
    h = ops.ConvT2dAct.apply(z, ws[0], bs[0], 1, 0, "relu", False, True, False)
    h = ops.ConvT2dAct.apply(h, ws[1], bs[1], 2, 1, "relu", False, True, True)
    h = ops.ConvT2dAct.apply(h, ws[2], bs[2], 2, 1, "relu", False, True, True)
    y = ops.ConvT2dAct.apply(h, ws[3], bs[3], 2, 1, "sigmoid", True, False, True)
    zd = z.detach().double().requires_grad_()
    wd = [w.detach().double().requires_grad_() for w in ws]
    bd = [b.detach().double().requires_grad_() for b in bs]
    r = torch.relu(F.conv_transpose2d(zd, wd[0], bd[0], stride=1, padding=0))
    r = torch.relu(F.conv_transpose2d(r, wd[1], bd[1], stride=2, padding=1))
    r = torch.relu(F.conv_transpose2d(r, wd[2], bd[2], stride=2, padding=1))
    yr = torch.sigmoid(F.conv_transpose2d(r, wd[3], bd[3], stride=2, padding=1))
    assert rel_err(y, yr) < 5e-6                     # two bf16x3 layers (convt_ts.cu) in the chain: 3-5e-6 each before the sigmoid
    gy = _rand(*yr.shape, seed=4)
    got = torch.autograd.grad(y, [z] + ws + bs, gy)
    ref = torch.autograd.grad(yr, [zd] + wd + bd, gy.double())
    for a, b in zip(got, ref):
        assert rel_err(a, b) < 3e-5


@pytest.mark.parametrize("dist,xform", [("laplace", "softmax"), ("normal", "exp_half"), ("normal", "exp_half_eps")])
@pytest.mark.parametrize("bk", [False, True])
def test_rsample(ops, dist, xform, bk):
    from oracle import synth
    import numpy as np
    M, K, B, L = 2, 7, 5, 20
    mu = _rand(M, B, L, seed=1).requires_grad_()
    raw = _rand(M, B, L, seed=2).requires_grad_()
    noise = synth.make_noise(dist, (M, K, B, L), np.random.RandomState(3)).cuda()
    nz = noise.permute(0, 2, 1, 3).contiguous() if bk else noise
    z, std = ops.RSample.apply(mu, raw, nz, dist, xform, bk)
    mud, rawd = mu.detach().double().requires_grad_(), raw.detach().double().requires_grad_()
    if xform == "softmax":
        sr = F.softmax(rawd, -1) * L + 1e-6
    else:
        sr = torch.exp(0.5 * rawd) + (1e-6 if xform == "exp_half_eps" else 0.0)
    zr = O.lat_rsample(dist, mud.unsqueeze(1), sr.unsqueeze(1), noise.double())
    zc = z.permute(0, 2, 1, 3) if bk else z
    assert rel_err(zc, zr) < 1e-6 and rel_err(std, sr) < 1e-6
    gz, gs = _rand(M, K, B, L, seed=5), _rand(M, B, L, seed=6)
    gzz = gz.permute(0, 2, 1, 3).contiguous() if bk else gz
    gmu, graw = torch.autograd.grad((z, std), (mu, raw), (gzz, gs))
    rmu, rraw = torch.autograd.grad((zr, sr), (mud, rawd), (gz.double(), gs.double()))
    assert rel_err(gmu, rmu) < 1e-5 and rel_err(graw, rraw) < 1e-5


@pytest.mark.parametrize("lik", ["normal", "laplace", "bernoulli"])
@pytest.mark.parametrize("D,x_div", [(784, 1), (3072, 1), (1024, 3), (12, 1)])
def test_loglik(ops, lik, D, x_div):
    B, Kk = 4, 3
    R = B * Kk
    recon = torch.sigmoid(_rand(R, D, seed=1)).requires_grad_()
    x = torch.rand(B, D, generator=torch.Generator().manual_seed(2)).cuda()
    if lik == "bernoulli":
        x = (x > 0.5).float()
        with torch.no_grad():
            recon[0, :4] = torch.tensor([0.0, 1.0, 1e-9, 1 - 1e-9])          # exercise the clamp
    c = 3.918
    lp, lo = ops.LogLik.apply(recon, x, x_div, lik, c, False)
    idx = ((torch.arange(R) // x_div) % B).cuda()
    if lik == "bernoulli":          # torch's clamp uses the fp32 eps: evaluate the reference in fp32 for this family
        rd = recon.detach().clone().requires_grad_()
        ref = (O.px_log_prob(lik, rd, x[idx]) * c).sum(-1)
    else:
        rd = recon.detach().double().requires_grad_()
        ref = (O.px_log_prob(lik, rd, x[idx].double()) * c).sum(-1)
    assert rel_err(lp, ref) < (1e-5 if lik == "bernoulli" else 2e-6)
    if lik != "bernoulli":          # (lp, lo) carries the fp64 row sum of the fp32 per-pixel terms
        assert rel_err(lp.detach().double() + lo.double(), ref) < 5e-8
        assert (lo.abs() <= lp.detach().abs() * 2.0 ** -23).all()
    g = _rand(R, seed=3)
    (gr,) = torch.autograd.grad(lp, recon, g)
    (rr,) = torch.autograd.grad(ref, rd, g.to(ref.dtype))
    assert rel_err(gr, rr) < 1e-5


def _moe_reference(z, mu, std, lpx, dist, mode, detach):
    M, K, B, L = z.shape
    ploc, pscale = torch.zeros(1, L, dtype=z.dtype, device=z.device), torch.ones(1, L, dtype=z.dtype, device=z.device)
    lws = []
    for r in range(M):
        lpz = O.lat_log_prob(dist, z[r], ploc, pscale).sum(-1)
        lq = torch.stack([O.lat_log_prob(dist, z[r], (mu[e].detach() if detach else mu[e]).unsqueeze(0),
                                         (std[e].detach() if detach else std[e]).unsqueeze(0)).sum(-1) for e in range(M)])
        lws.append(lpz + lpx[r].sum(0) - (torch.logsumexp(lq, 0) - math.log(M)))
    if mode == "iwae":
        lw = torch.cat(lws)
        return (torch.logsumexp(lw, 0) - math.log(lw.shape[0])).sum(), torch.stack(lws)
    if mode == "iwae_looser":
        lw = torch.stack(lws)
        return (torch.logsumexp(lw, 1) - math.log(K)).mean(0).sum(), lw
    if mode == "dreg":
        lw = torch.cat(lws)
        with torch.no_grad():
            w = (lw - torch.logsumexp(lw, 0, keepdim=True)).exp()
        return (w * lw).sum(), torch.stack(lws)
    lw = torch.stack(lws)
    with torch.no_grad():
        w = (lw - torch.logsumexp(lw, 1, keepdim=True)).exp()
    return (w * lw).mean(0).sum(), lw


@pytest.mark.parametrize("dist", ["laplace", "normal"])
@pytest.mark.parametrize("mode", ["iwae", "iwae_looser", "dreg", "dreg_looser"])
@pytest.mark.parametrize("M,K", [(2, 30), (3, 5), (2, 1), (2, 200)])
@pytest.mark.parametrize("with_lo", [False, True])
def test_moe_logw(ops, dist, mode, M, K, with_lo):
    B, L = 6, 20
    z = _rand(M, K, B, L, seed=1).requires_grad_()
    mu = _rand(M, B, L, seed=2, scale=0.5).requires_grad_()
    std = (torch.rand(M, B, L, generator=torch.Generator().manual_seed(3)) + 0.5).cuda().requires_grad_()
    lpx64 = _rand(M, M, K, B, seed=4).double() * 3 - (7000 if with_lo else 700)
    if with_lo:                     # (hi, lo) pair of an fp64 log-likelihood, as mmvae_loglik_fwd hands it over
        lpx64 = lpx64 + torch.rand(lpx64.shape, dtype=torch.float64, generator=torch.Generator().manual_seed(5)).cuda() * 1e-3
    lpx = lpx64.float().requires_grad_()
    lo = (lpx64 - lpx.detach().double()).float() if with_lo else None
    pl, ps = torch.zeros(L).cuda(), torch.ones(L).cuda()
    detach = mode.startswith("dreg")
    loss, lw, wt, per = ops.MoeLogW.apply(z, mu.detach() if detach else mu, std.detach() if detach else std, lpx, lo,
                                          pl, ps, dist, mode)
    zd, mud, stdd = (t.detach().double().requires_grad_() for t in (z, mu, std))
    lpxd = lpx64.clone().requires_grad_() if with_lo else lpx.detach().double().requires_grad_()
    ref, lwr = _moe_reference(zd, mud, stdd, lpxd, dist, mode, detach)
    assert rel_err(lw, lwr) < 1e-6
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())          # vs fp64
    assert abs(per.sum().item() - loss.item()) <= 1e-5 * abs(ref.item())
    ins = (z, lpx) if detach else (z, mu, std, lpx)
    insd = (zd, lpxd) if detach else (zd, mud, stdd, lpxd)
    g = torch.autograd.grad(loss * 1.7, ins)
    gr = torch.autograd.grad(ref * 1.7, insd)
    for a, b in zip(g, gr):
        assert rel_err(a, b) < 1e-5          # lw and the softmax are formed in fp64: no |lw|*eps noise on the weights


@pytest.mark.parametrize("dist", ["laplace", "normal"])
def test_elbo(ops, dist):
    M, K, B, L = 2, 4, 5, 20
    z = _rand(M, K, B, L, seed=1, scale=0.3)
    mu = _rand(M, B, L, seed=2, scale=0.3).requires_grad_()
    std = (torch.rand(M, B, L, generator=torch.Generator().manual_seed(3)) * 0.2 + 0.9).cuda().requires_grad_()
    lpx = (_rand(M, K, B, seed=4) - 50).requires_grad_()
    pl, ps = torch.zeros(L).cuda(), torch.ones(L).cuda()
    loss, terms, per = ops.Elbo.apply(z, mu, std, lpx, pl, ps, dist)
    mud, stdd, lpxd, zd = (t.detach().double().requires_grad_() for t in (mu, std, lpx, z))
    ploc, pscale = torch.zeros(1, L, dtype=torch.double).cuda(), torch.ones(1, L, dtype=torch.double).cuda()
    tot = 0
    for r in range(M):
        tot = tot - O.lat_kl(dist, mud[r], stdd[r], ploc, pscale).sum(-1).unsqueeze(0)
        for d in range(M):
            if d == r:
                w = 1.0
            else:
                w = (O.lat_log_prob(dist, zd[d], mud[r].unsqueeze(0), stdd[r].unsqueeze(0))
                     - O.lat_log_prob(dist, zd[d], mud[d].unsqueeze(0), stdd[d].unsqueeze(0)).detach()).sum(-1).exp()
            tot = tot + w * lpxd[d]
    ref = (tot / M).mean(0).sum()
    assert abs(loss.item() - ref.item()) <= 2e-5 * abs(ref.item())
    g = torch.autograd.grad(loss, (mu, std, lpx))
    gr = torch.autograd.grad(ref, (mud, stdd, lpxd))
    for a, b in zip(g, gr):
        assert rel_err(a, b) < 1e-4


def test_lse_and_eval_logw(ops):
    B, K, L, E, D = 7, 250, 20, 2, 2
    a = _rand(B, K, seed=1) * 5 - 3000
    mx, se = ops.lse_partial(a[:, :100].contiguous())
    mx, se = ops.lse_partial(a[:, 100:].contiguous(), mx, se)
    ll = ops.lse_finish(mx, se, 300)
    ref = torch.logsumexp(a.double(), 1) - math.log(300)
    assert (ll.double() - ref).abs().max().item() < 1e-3
    z = _rand(B, K, L, seed=2)
    mu, std = _rand(E, B, L, seed=3), (torch.rand(E, B, L) + 0.5).cuda()
    lpx = _rand(D, B, K, seed=4) - 100
    pl, ps = torch.zeros(L).cuda(), torch.ones(L).cuda()
    for dist in ("laplace", "normal"):
        lpz = O.lat_log_prob(dist, z.double(), pl.double(), ps.double()).sum(-1)
        lq = torch.stack([O.lat_log_prob(dist, z.double(), mu[e].double().unsqueeze(1), std[e].double().unsqueeze(1)).sum(-1)
                          for e in range(E)])
        base = lpx.double().sum(0) + lpz
        for qmode, q in (("mixture_half", torch.logsumexp(lq, 0) / 2), ("mixture", torch.logsumexp(lq, 0) - math.log(E)),
                         ("single", lq[0]), ("none", 0 * lpz)):
            out = ops.eval_logw(z, mu, std, lpx, pl, ps, dist, qmode)
            assert (out.double() - (base - q)).abs().max().item() < 1e-3
        out = ops.eval_logw(z, None, None, lpx, pl, ps, dist, "external", lq_ext=lq[1].float().contiguous())
        assert (out.double() - (base - lq[1])).abs().max().item() < 1e-3


def test_adam_amsgrad_matches_torch(ops):
    p0 = _rand(1000, seed=1)
    p = p0.clone()
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref], lr=1e-3, amsgrad=True)
    m, v, vm = torch.zeros_like(p), torch.zeros_like(p), torch.zeros_like(p)
    step = torch.zeros(1).cuda()
    for i in range(5):
        g = _rand(1000, seed=10 + i)
        ref.grad = g.clone()
        opt.step()
        ops.adam_amsgrad_(p, g, m, v, vm, step, lr=1e-3)
    assert step.item() == 5
    assert rel_err(p, ref) < 1e-6


def test_cpu_tensor_is_rejected(ops):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.LinearAct.apply(torch.zeros(2, 3), torch.zeros(4, 3).cuda(), None, None)


@pytest.mark.parametrize("L,flow,R", [(20, "MAF", 1), (20, "MAF", 63), (20, "MAF", 5000), (20, "IAF", 65), (20, "IAF", 4097),
                                     (30, "MAF", 777), (17, "IAF", 300)])
def test_fused_sequential_made_block_matches_gemm_formulation(ops, L, flow, R):
    """mmvae_made_seq_fwd (one kernel per flow block, every hidden unit evaluated once) against the L-pass formulation
    (4 masked GEMMs + one affine step per latent column) it replaces, and against a float64 torch restatement of
    maf_model.py:91-104 / iaf_model.py:60-78.  Tolerance: 2e-5 relative on outputs of magnitude ~1 (fp32 summation order)."""
    from mmvae_b200 import flows
    torch.manual_seed(L + R)
    f = getattr(flows, flow)(L, 2, 3, 128).cuda()
    with torch.no_grad():
        for p in f.parameters():
            p.mul_(1.7)                                    # livelier log-variances than the default init
    x = torch.randn(R, L, device="cuda")
    assert all(ops.made_seq_plan(m).ok for m in f.net)
    call = f.inverse if flow == "MAF" else f.forward
    with torch.no_grad():
        n0 = ops.N_CALLS
        o_f = call(x)
        fused_calls = ops.N_CALLS - n0
        flows.FUSED_SEQUENTIAL = False
        try:
            n0 = ops.N_CALLS
            o_g = call(x)
            gemm_calls = ops.N_CALLS - n0
        finally:
            flows.FUSED_SEQUENTIAL = True
    assert fused_calls * 10 < gemm_calls
    assert rel_err(o_f.out, o_g.out) < 2e-5 and rel_err(o_f.log_abs_det_jac, o_g.log_abs_det_jac) < 2e-5
    # float64 restatement
    xs = x.double().cpu()
    blocks = list(f.net)[::-1] if flow == "MAF" else list(f.net)
    ladj = torch.zeros(R, dtype=torch.float64)
    y = xs
    for made in blocks:
        if flow == "MAF":
            y = y.flip(1)
        ws = [(l.weight * l.mask).double().cpu() for l in made.layers()]
        bs = [l.bias.double().cpu() for l in made.layers()]
        cur = torch.zeros_like(y)
        for i in range(L):
            h = cur
            for j, (w, b) in enumerate(zip(ws, bs)):
                h = h @ w.t() + b
                if j < 3:
                    h = torch.relu(h)
            mu, lv = h[:, :L], h[:, L:]
            if flow == "MAF":
                cur[:, i] = y[:, i] * torch.exp(lv[:, i]) + mu[:, i]
                ladj = ladj + lv[:, i]
            else:
                cur[:, i] = (y[:, i] - mu[:, i]) * torch.exp(-lv[:, i])
                ladj = ladj - lv[:, i]
        y = cur if flow == "MAF" else cur.flip(1)
    assert rel_err(o_f.out, y) < 1e-5 and rel_err(o_f.log_abs_det_jac, ladj) < 1e-5
    # with gradients required the differentiable formulation runs
    xg = x.clone().requires_grad_(True)
    o = call(xg)
    o.out.sum().backward()
    assert xg.grad is not None


def test_fused_sequential_made_rejects_unsupported_shapes(ops):
    from mmvae_b200 import flows
    assert not ops.made_seq_plan(flows.MADE(2, [128] * 3)).ok          # one degree, 128 units in its group
    assert not ops.made_seq_plan(flows.MADE(20, [128] * 2)).ok         # two hidden layers
    m = flows.MADE(20, [128] * 3)
    m.net[0].mask.zero_()
    assert not ops.made_seq_plan(m).ok                                  # masks that are not the sequential ones


@pytest.mark.parametrize("R", [1, 16, 17, 128, 300, 1024])
@pytest.mark.parametrize("L,H,n_hidden", [(20, 128, 3), (8, 64, 1), (12, 32, 2)])
def test_made_fused_training_kernels(ops, R, L, H, n_hidden):
    """csrc/made_train.cu (one forward kernel, one data-gradient kernel, one weight-gradient kernel per MADE) against the
    per-layer formulation of the same module (mask, GEMM, activation backward, column sums) and against fp64 torch: output,
    gradient w.r.t. the input and every masked weight / bias gradient; ragged row counts; bit-identical on a second pass."""
    from mmvae_b200 import flows
    torch.manual_seed(R + L)
    made = flows.MADE(L, [H] * n_hidden).cuda()
    if R == 17:                                                   # parameters as views at odd offsets of one flat buffer (train.Trainer)
        ps = list(made.parameters())
        flat = torch.empty(sum(p.numel() + 1 for p in ps) + 1, device="cuda")
        off = 1
        for p in ps:
            flat[off:off + p.numel()].copy_(p.data.reshape(-1))
            p.data = flat[off:off + p.numel()].view_as(p)
            off += p.numel() + 1
    x = _rand(R, L, seed=5).requires_grad_()
    gy = _rand(R, 2 * L, seed=6)
    params = [p for p in made.parameters()]

    def run(fused):
        ops.MADE_FUSED = fused
        try:
            y = made.raw(x)
            grads = torch.autograd.grad(y, [x] + params, gy)
        finally:
            ops.MADE_FUSED = True
        return y.detach(), grads
    n0 = ops.N_CALLS
    y1, g1 = run(True)
    assert ops.N_CALLS - n0 == 3                                   # forward, data gradient, weight gradients
    y2, g2 = run(False)
    y3, g3 = run(True)
    assert torch.equal(y1, y3) and all(torch.equal(a, b) for a, b in zip(g1, g3))
    # fp64 reference of the module
    h = x.detach().double()
    h.requires_grad_()
    lay = made.layers()
    ws = [l.weight.detach().double().requires_grad_() for l in lay]
    bs = [l.bias.detach().double().requires_grad_() for l in lay]
    z = h
    for i, l in enumerate(lay):
        z = torch.nn.functional.linear(z, ws[i] * l.mask.double(), bs[i])
        if i < len(lay) - 1:
            z = z.relu()
    ref = torch.autograd.grad(z, [h] + [t for wb in zip(ws, bs) for t in wb], gy.double())
    assert rel_err(y1, z) < 2e-6 and rel_err(y2, z) < 2e-6
    for a, b, r in zip(g1, g2, ref):
        if r.abs().max().item() == 0.0:
            assert a.abs().max().item() == 0.0
            continue
        assert rel_err(a, r) < 5e-6, (a.shape, rel_err(a, r))
        assert rel_err(b, r) < 5e-6
