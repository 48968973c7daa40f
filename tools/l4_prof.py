"""Runs the three direct last-layer kernels (Cin=32 -> Cout=3, 16x16 -> 32x32) a few times: target for ncu captures."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib
lib = _lib.load()
p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
N, H, Cin, Cout = 7680, 16, 32, 3
x = torch.randn(N, H, H, Cin, device="cuda").relu_()
w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
b = torch.randn(Cout, device="cuda")
y = torch.empty(N, Cout, 2 * H, 2 * H, device="cuda")
g = torch.randn(N, Cout, 2 * H, 2 * H, device="cuda")
gx = torch.empty_like(x); gw = torch.empty_like(w); gb = torch.empty_like(b)
ws = torch.empty(lib.mmvae_convt_small_wgrad_workspace_bytes(N, Cin, Cout) // 4, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    _lib.check(lib.mmvae_convt_small_fwd(p(x), p(w), p(b), p(y), N, H, H, Cin, Cout, 2, st()))
    _lib.check(lib.mmvae_convt_small_dgrad(p(g), p(y), p(w), p(x), p(gx), N, H, H, Cin, Cout, 2, st()))
    _lib.check(lib.mmvae_convt_small_wgrad(p(x), p(g), p(y), p(gw), p(gb), p(ws), N, H, H, Cin, Cout, 2, st()))
torch.cuda.synchronize()
