"""GPU check + timing of the second-generation transposed-convolution kernels (convt_ts.cu) against fp64 and the first
generation (gemm_tc.cu).  python tools/ts_conv_check.py [N]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from mmvae_b200 import ops


def rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def case(N, cin, cout, H, check=True, relu_in=True):
    g = torch.Generator().manual_seed(N + cin)
    x = torch.randn(N, H, H, cin, generator=g).cuda()
    if relu_in:
        x = x.relu()
    x = x.permute(0, 3, 1, 2).requires_grad_()
    w = (0.1 * torch.randn(cin, cout, 4, 4, generator=g)).cuda().requires_grad_()
    b = torch.randn(cout, generator=g).cuda().requires_grad_()
    gy = torch.randn(N, 2 * H, 2 * H, cout, generator=g).cuda().permute(0, 3, 1, 2)
    out = {}
    for ts in (True, False):
        ops.ALLOW_TS = ts
        y = ops.ConvT2dAct.apply(x, w, b, 2, 1, "relu", False, True, relu_in)     # preact_grad: gy is w.r.t. the pre-activation
        gx, gw, gb = torch.autograd.grad(y, (x, w, b), gy)
        torch.cuda.synchronize()
        out[ts] = (y.detach(), gx, gw, gb)
    if check:
        xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
        pre = F.conv_transpose2d(xd, wd, bd, stride=2, padding=1)
        yr = pre.relu()
        rx, rw, rb = torch.autograd.grad(pre, (xd, wd, bd), gy.double())
        if relu_in:
            rx = rx * (xd > 0)
        for ts in (True, False):
            y, gx, gw, gb = out[ts]
            print(f"  N={N} {cin}->{cout} H={H} {'TS ' if ts else 'old'}: y {rel(y, yr):.2e}  gx {rel(gx, rx):.2e}  gw {rel(gw, rw):.2e}  gb {rel(gb, rb):.2e}  "
                  f"max|y-yr| {(y.double() - yr).abs().max().item():.2e}", flush=True)
    for ts in (True, False):
        ops.ALLOW_TS = ts
        with torch.no_grad():
            tf = timeit(lambda: ops.ConvT2dAct.apply(x, w, b, 2, 1, "relu", False, True, relu_in))
        y = ops.ConvT2dAct.apply(x, w, b, 2, 1, "relu", False, True, relu_in)
        tb = timeit(lambda: torch.autograd.grad(y, (x,), gy, retain_graph=True))
        print(f"  N={N} {cin}->{cout} H={H} {'TS ' if ts else 'old'}: fwd (incl. weight pack) {tf:.1f} us, dgrad {tb:.1f} us", flush=True)
    ops.ALLOW_TS = True


if __name__ == "__main__":
    big = int(sys.argv[1]) if len(sys.argv) > 1 else 7680
    for N in (2, 5, 16, 33):
        case(N, 64, 32, 8)
        case(N, 128, 64, 4)
    case(big, 64, 32, 8, check=False)
    case(big, 128, 64, 4, check=False)
    case(300, 64, 32, 8)
    case(300, 128, 64, 4)
