"""Ablation timing of convt_fwd9_kernel (results are wrong by construction; only the times mean anything).
bits: 1 no global stores, 2 no A loads, 4 no B loads, 8 no hi/lo conversion, 16 no MMAs (first only)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib
lib = _lib.load()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def bench(fn, n=10):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
N = 7680
for (H, Cin, Cout) in ((4, 128, 64), (8, 64, 32)):
    x = torch.randn(N, H, H, Cin, device="cuda"); w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
    b = torch.randn(Cout, device="cuda"); wf = torch.empty(16 * Cin * Cout, device="cuda")
    lib.mmvae_convt_weight_classes(p(w), p(wf), Cin, Cout, st())
    y = torch.empty(N, 2 * H, 2 * H, Cout, device="cuda")
    for mode in (0, 1):
        for dbg in ([int(a) for a in sys.argv[1:]] or (0, 1, 2, 4, 6, 8, 16, 1 | 8, 2 | 4 | 8, 1 | 2 | 4 | 8, 1 | 2 | 4 | 8 | 16, 2 | 4 | 8 | 16, 1 | 8 | 16)):
            lib.mmvae_gemm_tc_set_mode(mode | (dbg << 8))
            t = bench(lambda: _lib.check(lib.mmvae_convt_tc_fwd(p(x), p(wf), p(b), p(y), N, H, H, Cin, Cout, 1, st())))
            print(f"H={H} Cin={Cin} Cout={Cout} passes={3 if mode == 0 else 1} dbg={dbg:2d} "
                  f"[{'nostore ' if dbg & 1 else ''}{'noA ' if dbg & 2 else ''}{'noB ' if dbg & 4 else ''}{'noconv ' if dbg & 8 else ''}{'nomma' if dbg & 16 else ''}]: {t:7.1f} us", flush=True)
lib.mmvae_gemm_tc_set_mode(0)
