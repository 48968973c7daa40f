"""Where the clocks of one CTA of convt_ts_kernel go (TRACE instantiation: clock64 around every wait of every role of CTA 0).
python tools/ts_trace.py [bits]      bits = ablation bits of tools/ts_time.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mmvae_b200 import _lib
lib = _lib.load()
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
bits = int(sys.argv[1]) if len(sys.argv) > 1 else 0
NAMES = {0: ["wait empty", "", "", "", "", "", "", "total"],
         2: ["wait tempty", "wait a_full", "wait b_full", "issue+commit", "", "", "", "total"],
         3: ["wait full", "split", "bar.sync", "wait done", "copy+st+arrive", "fence+release", "", "total"],
         4: ["wait tfull", "", "", "", "", "", "", "total"]}
ROLE = {0: "producer", 2: "issuer", 3: "converter warp 3", 4: "epilogue warp 11"}
N = 7680
for cin, cout, H in ((64, 32, 8), (128, 64, 4)):
    x = torch.randn(N, H, H, cin, device="cuda").relu()
    w = 0.1 * torch.randn(cin, cout, 4, 4, device="cuda")
    b = torch.randn(cout, device="cuda")
    g = torch.randn(N, 2 * H, 2 * H, cout, device="cuda")
    y = torch.empty(N, 2 * H, 2 * H, cout, device="cuda")
    gx = torch.empty_like(x)
    ne = lib.mmvae_convt_ts_packed_elems(cin, cout)
    wf = torch.empty(ne, device="cuda", dtype=torch.int16)
    wd = torch.empty(ne, device="cuda", dtype=torch.int16)
    _lib.check(lib.mmvae_convt_ts_pack_weights(P(w), P(wf), P(wd), cin, cout, None))
    tr = torch.zeros(64 + 2 * 148, device="cuda", dtype=torch.int64)
    lib.mmvae_gemm_tc_set_mode(bits << 8)
    for name, fn in (("fwd", lambda: lib.mmvae_convt_ts_fwd(P(x), P(wf), P(b), P(y), N, H, H, cin, cout, 1, None)),
                     ("dgrad", lambda: lib.mmvae_convt_ts_dgrad(P(g), P(wd), P(x), P(gx), N, H, H, cin, cout, None))):
        fn(); torch.cuda.synchronize()
        lib.mmvae_convt_ts_set_trace(ctypes.c_void_p(tr.data_ptr()))
        tr.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        _lib.check(fn()); torch.cuda.synchronize()
        e0.record(); _lib.check(fn()); e1.record(); torch.cuda.synchronize()
        lib.mmvae_convt_ts_set_trace(None)
        t = tr.cpu().tolist()
        st, en = t[64::2], t[65::2]
        t0 = min(st)
        dur = sorted(e - s for s, e in zip(st, en))
        print(f"   traced kernel {e0.elapsed_time(e1) * 1e3:.1f} us by events; CTA starts {min(st) - t0}..{max(st) - t0} ns, issuer ends "
              f"{min(en) - t0}..{max(en) - t0} ns; per-CTA busy ns min/median/max {dur[0]}/{dur[len(dur) // 2]}/{dur[-1]}")
        tiles = (N * H * H // 128 + 147) // 148
        print(f"{cin}->{cout} H={H} {name} (dbg {bits}), CTA 0, {tiles} tiles: clocks per tile")
        for r in (0, 2, 3, 4):
            vals = t[r * 8:(r + 1) * 8]
            print(f"   {ROLE[r]:18s} " + "  ".join(f"{n} {v / tiles:7.0f}" for n, v in zip(NAMES[r], vals) if n))
    lib.mmvae_gemm_tc_set_mode(0)
