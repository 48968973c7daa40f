"""One launch of each convt_ts kernel at C2 size (for ncu).  python tools/ts_one.py [l3|l2]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mmvae_b200 import _lib
lib = _lib.load()
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
which = sys.argv[1] if len(sys.argv) > 1 else "l3"
N, (cin, cout, H) = 7680, ((64, 32, 8) if which == "l3" else (128, 64, 4))
x = torch.randn(N, H, H, cin, device="cuda").relu()
w = 0.1 * torch.randn(cin, cout, 4, 4, device="cuda")
b = torch.randn(cout, device="cuda")
g = torch.randn(N, 2 * H, 2 * H, cout, device="cuda")
y = torch.empty(N, 2 * H, 2 * H, cout, device="cuda")
gx = torch.empty_like(x)
ne = lib.mmvae_convt_ts_packed_elems(cin, cout)
wf = torch.empty(ne, device="cuda", dtype=torch.int16)
wd = torch.empty(ne, device="cuda", dtype=torch.int16)
nb = lib.mmvae_convt_ts_wgrad_workspace_bytes(N, H, H, cin, cout)
ws = torch.empty(nb // 4, device="cuda")
gw = torch.empty_like(w)
gb = torch.empty_like(b)
for _ in range(2):
    _lib.check(lib.mmvae_convt_ts_pack_weights(P(w), P(wf), P(wd), cin, cout, None))
    _lib.check(lib.mmvae_convt_ts_fwd(P(x), P(wf), P(b), P(y), N, H, H, cin, cout, 1, None))
    _lib.check(lib.mmvae_convt_ts_dgrad(P(g), P(wd), P(x), P(gx), N, H, H, cin, cout, None))
    _lib.check(lib.mmvae_convt_ts_wgrad(P(x), P(g), P(gw), P(gb), P(ws), nb, N, H, H, cin, cout, None))
torch.cuda.synchronize()
print("ok")
