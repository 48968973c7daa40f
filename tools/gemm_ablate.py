"""Ablation timing of gemm_tc_kernel in its dgrad / wgrad convolution modes (results wrong by construction).
bits: 1 no epilogue after the TMEM read, 2|4 no operand loads, 8 no hi/lo conversion, 16 no MMAs (first only)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops
lib = _lib.load()
p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def bench(fn, n=10):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
N = 7680
names = {1: "noepi", 6: "noload", 8: "noconv", 16: "nomma", 8192: "3 UMMAs per k-step (no hi*[hi|lo] merge)"}
for (H, Cin, Cout) in ((4, 128, 64), (8, 64, 32)):
    x = torch.randn(N, H, H, Cin, device="cuda"); w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
    g = torch.randn(N, 2 * H, 2 * H, Cout, device="cuda")
    wp = ops.permute_weight(w); gx = torch.empty_like(x); gwp = torch.empty(Cin, 16 * Cout, device="cuda")
    nb = lib.mmvae_convt_tc_wgrad_workspace_bytes(N, H, H, Cin, Cout); ws = torch.empty(max(nb // 4, 1), device="cuda")
    for mode in (0, 1):
        for dbg in ([int(a) for a in sys.argv[1:]] or (0, 8192, 1, 6, 8, 16, 1 | 8, 1 | 6 | 8, 1 | 6 | 8 | 16)):
            lib.mmvae_gemm_tc_set_mode(mode | (dbg << 8))
            td = bench(lambda: _lib.check(lib.mmvae_convt_tc_dgrad(p(g), p(wp), p(x), p(gx), N, H, H, Cin, Cout, st())))
            tw = bench(lambda: _lib.check(lib.mmvae_convt_tc_wgrad(p(x), p(g), p(gwp), p(ws), nb, N, H, H, Cin, Cout, st())))
            tag = " ".join(v for k, v in names.items() if dbg & k)
            print(f"H={H} Cin={Cin} Cout={Cout} passes={3 if mode == 0 else 1} dbg={dbg:4d} [{tag}]: dgrad {td:7.1f} us  wgrad {tw:7.1f} us", flush=True)
lib.mmvae_gemm_tc_set_mode(0)
