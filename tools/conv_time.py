"""Times the implicit-GEMM convT kernels (fwd / dgrad / wgrad) at the C2 decoder shapes, in 3xTF32 mode and in the
single-pass TF32 probe mode (separates MMA-bound from load/latency-bound).   python tools/conv_time.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops

lib = _lib.load()
p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def bench(fn, n=10):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


N = int(sys.argv[1]) if len(sys.argv) > 1 else 7680
for (H, Cin, Cout) in ((4, 128, 64), (8, 64, 32)):
    x = torch.randn(N, H, H, Cin, device="cuda")
    w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
    b = torch.randn(Cout, device="cuda")
    g = torch.randn(N, 2 * H, 2 * H, Cout, device="cuda")
    wf = torch.empty(16 * Cin * Cout, device="cuda")
    lib.mmvae_convt_weight_classes(p(w), p(wf), Cin, Cout, st())
    wp = ops.permute_weight(w)
    y = torch.empty(N, 2 * H, 2 * H, Cout, device="cuda")
    gx = torch.empty_like(x)
    gwp = torch.empty(Cin, 16 * Cout, device="cuda")
    nb = lib.mmvae_convt_tc_wgrad_workspace_bytes(N, H, H, Cin, Cout)
    ws = torch.empty(max(nb // 4, 1), device="cuda")
    flop = 2.0 * N * H * H * 16 * Cin * Cout
    for mode in ((1,) if os.environ.get("PROBE") else (0, 1)):
        lib.mmvae_gemm_tc_set_mode(mode)
        t_f = bench(lambda: _lib.check(lib.mmvae_convt_tc_fwd(p(x), p(wf), p(b), p(y), N, H, H, Cin, Cout, 1, st())))
        t_d = bench(lambda: _lib.check(lib.mmvae_convt_tc_dgrad(p(g), p(wp), p(x), p(gx), N, H, H, Cin, Cout, st())))
        t_w = bench(lambda: _lib.check(lib.mmvae_convt_tc_wgrad(p(x), p(g), p(gwp), p(ws), nb, N, H, H, Cin, Cout, st())))
        print(f"H={H} Cin={Cin} Cout={Cout} mode={'3xTF32' if mode == 0 else '1xTF32'}: "
              f"fwd {t_f:7.1f} us ({flop / t_f / 1e6:6.1f} TF)  dgrad {t_d:7.1f} us ({flop / t_d / 1e6:6.1f} TF)  "
              f"wgrad {t_w:7.1f} us ({flop / t_w / 1e6:6.1f} TF)", flush=True)
lib.mmvae_gemm_tc_set_mode(0)

# last decoder layer (Cin=32 -> Cout=3, 16x16 -> 32x32, Sigmoid, NCHW out): the direct SIMT kernels
H, Cin, Cout = 16, 32, 3
x = torch.randn(N, H, H, Cin, device="cuda").relu_()
w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
b = torch.randn(Cout, device="cuda")
y = torch.empty(N, Cout, 2 * H, 2 * H, device="cuda")
g = torch.randn(N, Cout, 2 * H, 2 * H, device="cuda")
gx = torch.empty_like(x)
gw = torch.empty_like(w)
gb = torch.empty_like(b)
ws = torch.empty(lib.mmvae_convt_small_wgrad_workspace_bytes(N, Cin, Cout) // 4, device="cuda")
t_f = bench(lambda: _lib.check(lib.mmvae_convt_small_fwd(p(x), p(w), p(b), p(y), N, H, H, Cin, Cout, 2, st())))
t_d = bench(lambda: _lib.check(lib.mmvae_convt_small_dgrad(p(g), p(y), p(w), p(x), p(gx), N, H, H, Cin, Cout, 2, st())))
t_w = bench(lambda: _lib.check(lib.mmvae_convt_small_wgrad(p(x), p(g), p(y), p(gw), p(gb), p(ws), N, H, H, Cin, Cout, 2, st())))
flop = 2.0 * N * H * H * 16 * Cin * Cout
print(f"L4 direct (H=16 Cin=32 Cout=3): fwd {t_f:7.1f} us ({flop / t_f / 1e6:5.1f} TF)  dgrad {t_d:7.1f} us ({flop / t_d / 1e6:5.1f} TF)  "
      f"wgrad {t_w:7.1f} us ({flop / t_w / 1e6:5.1f} TF)", flush=True)
