"""Direct C-ABI timing of convt_ts.cu (forward / dgrad) at C2 size with ablation bits.  python tools/ts_time.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mmvae_b200 import _lib

lib = _lib.load()
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None


def timeit(fn, n=20):
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


def run(N, cin, cout, H, bits_list):
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
    _lib.check(lib.mmvae_convt_ts_pack_weights(P(w), P(wf), P(wd), cin, cout, None))
    tp = timeit(lambda: lib.mmvae_convt_ts_pack_weights(P(w), P(wf), P(wd), cin, cout, None))
    print(f"{cin}->{cout} H={H} N={N}: pack {tp:.1f} us")
    for bits in bits_list:
        lib.mmvae_gemm_tc_set_mode(bits << 8)
        tf = timeit(lambda: _lib.check(lib.mmvae_convt_ts_fwd(P(x), P(wf), P(b), P(y), N, H, H, cin, cout, 1, None)))
        td = timeit(lambda: _lib.check(lib.mmvae_convt_ts_dgrad(P(g), P(wd), P(x), P(gx), N, H, H, cin, cout, None)))
        tw = timeit(lambda: _lib.check(lib.mmvae_convt_ts_wgrad(P(x), P(g), P(gw), P(gb), P(ws), nb, N, H, H, cin, cout, None)))
        lib.mmvae_gemm_tc_set_mode(0)
        print(f"  dbg {bits:2d}: fwd {tf:7.1f} us   dgrad {td:7.1f} us   wgrad+bias {tw:7.1f} us", flush=True)


if __name__ == "__main__":
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 7680
    bl = [0, 2, 4, 8, 2 | 4 | 8] if len(sys.argv) < 3 else [0, 1, 16, 14, 15, 30, 31, 1 | 16, 8 | 16, 2 | 4]
    run(N, 64, 32, 8, bl)
    run(N, 128, 64, 4, bl)
