"""Times the GEMM shapes of the C2 step whose K or N is the latent size (20) on the tcgen05 path (inner extent below the
32-float TMA box, zero-filled) against the exact-fp32 SIMT kernel, and the K=1000 evaluator leg both ways."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops
lib = _lib.load()


def bench(fn, n=20):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


SHAPES = [  # M, N, K, a_k, b_k, what
    (7680, 512, 20, 1, 1, "MNIST dec L0 fwd"), (7680, 20, 512, 1, 0, "MNIST dec L0 dgrad"), (512, 20, 7680, 0, 0, "MNIST dec L0 wgrad"),
    (7680, 2048, 20, 1, 0, "SVHN dec L1 fwd"), (7680, 20, 2048, 1, 1, "SVHN dec L1 dgrad"), (20, 2048, 7680, 0, 0, "SVHN dec L1 wgrad"),
    (128, 20, 512, 1, 1, "enc head fwd"), (128, 512, 20, 1, 0, "enc head dgrad"), (20, 512, 128, 0, 0, "enc head wgrad"),
    (100000, 512, 20, 1, 1, "eval MNIST dec L0"), (32000, 2048, 20, 1, 0, "eval SVHN dec L1"),
]
for M, N, K, ak, bk, what in SHAPES:
    A = torch.randn((M, K) if ak else (K, M), device="cuda")
    B = torch.randn((N, K) if bk else (K, N), device="cuda")
    out = torch.empty(M, N, device="cuda")
    ts = []
    for dbg in (128, 0, 256):
        lib.mmvae_gemm_tc_set_mode(dbg << 8)
        ts.append(bench(lambda: ops.gemm(A, B, None, M, N, K, ak, bk, out=out)))
    print(f"{what:22s} M={M:6d} N={N:5d} K={K:5d} a_k={ak} b_k={bk}:  strict gate {ts[0]:7.1f} us   thin M|N on tcgen05 {ts[1]:7.1f} us   + K<32 on tcgen05 {ts[2]:7.1f} us", flush=True)
lib.mmvae_gemm_tc_set_mode(0)
if "--eval" in sys.argv:
    sys.argv = sys.argv[:1]
    import bench as Bn
    dev = torch.device("cuda", 0)
    model = Bn.make_model(dev)
    for dbg in (128, 0, 256):
        lib.mmvae_gemm_tc_set_mode(dbg << 8)
        ev = Bn.run_eval(model, dev, 64, Bn.peaks())
        tag = {128: 'strict', 0: 'thin M|N on tc', 256: '+ K<32 on tc'}[dbg]
        print(f"eval leg, {tag} gate: {ev['value']:.0f} datapoints/s ({ev['s_per_pass'] * 1e3:.2f} ms/pass)", flush=True)
    lib.mmvae_gemm_tc_set_mode(0)
