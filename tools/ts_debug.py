import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from mmvae_b200 import _lib
lib = _lib.load()
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
for (cin, cout, H) in ((64, 32, 8), (128, 64, 4)):
  for N in (6, 33, 39, 76, 77, 78, 149, 150, 299):
    gen = torch.Generator().manual_seed(N)
    w = (0.1 * torch.randn(cin, cout, 4, 4, generator=gen)).cuda()
    g = torch.randn(N, 2 * H, 2 * H, cout, generator=gen).cuda()
    x = torch.randn(N, H, H, cin, generator=gen).cuda()
    wd_ = w.double().requires_grad_()
    pre = F.conv_transpose2d(x.double().permute(0, 3, 1, 2), wd_, None, stride=2, padding=1)
    rw, = torch.autograd.grad(pre, wd_, g.double().permute(0, 3, 1, 2))
    rb = g.double().sum((0, 1, 2))
    nb = lib.mmvae_convt_ts_wgrad_workspace_bytes(N, H, H, cin, cout)
    outs = []
    for rep in range(3):
        ws = torch.full((nb // 4,), float("nan"), device="cuda")
        gw = torch.full_like(w, float("nan")); gb = torch.full((cout,), float("nan"), device="cuda")
        _lib.check(lib.mmvae_convt_ts_wgrad(P(x), P(g), P(gw), P(gb), P(ws), nb, N, H, H, cin, cout, None))
        torch.cuda.synchronize()
        outs.append((gw.clone(), gb.clone()))
    gw, gb = outs[0]
    err = (gw.double() - rw).abs()
    print(f"{cin}->{cout} N={N}: gw rel {((gw.double()-rw).norm()/rw.norm()).item():.2e} gb rel {((gb.double()-rb).norm()/rb.norm()).item():.2e} "
          f"deterministic {all(torch.equal(outs[0][0], o[0]) for o in outs)} nan {torch.isnan(gw).sum().item()} max {err.max().item():.3e}")
    if err.max().item() > 1e-2 * rw.abs().max().item():
        idx = (err > 1e-2 * rw.abs().max().item()).nonzero()
        print("   bad count", idx.shape[0], "ci", sorted(set(idx[:, 0].tolist()))[:8], "co", sorted(set(idx[:, 1].tolist()))[:8], "ky", sorted(set(idx[:, 2].tolist())), "kx", sorted(set(idx[:, 3].tolist())))
