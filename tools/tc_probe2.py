"""Accumulation-chain probe: error of the 3xTF32 tcgen05 GEMM vs the SIMT fp32 GEMM as K grows (no split-K), and timing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200
from mmvae_b200 import ops, _lib
lib = _lib.load()
torch.manual_seed(0)

def bench(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for (M, N, K, a_k, b_k) in [(2048, 2048, 128, 1, 1), (2048, 2048, 512, 1, 1), (2048, 2048, 2048, 1, 1), (2048, 2048, 8192, 1, 1),
                            (122880, 1024, 128, 1, 1), (122880, 128, 1024, 1, 1), (128, 1024, 122880, 0, 0),
                            (491520, 512, 64, 1, 1), (491520, 64, 512, 1, 1), (64, 512, 491520, 0, 0),
                            (1966080, 48, 32, 1, 1), (1966080, 32, 48, 1, 1), (32, 48, 1966080, 0, 0),
                            (7680, 784, 512, 1, 1), (7680, 512, 784, 1, 1), (784, 512, 7680, 0, 0)]:
    A = torch.rand(M, K, device="cuda") - 0.3
    B = torch.rand(N, K, device="cuda") - 0.3
    a_s = A if a_k else A.t().contiguous()
    b_s = B if b_k else B.t().contiguous()
    ref = (A.double() @ B.double().t()) if M * N * K < 3e11 else None
    res = {}
    for name, tc, mode in (("simt", False, 0), ("tc3", True, 0), ("tc1", True, 1)):
        ops.ALLOW_TC = tc
        lib.mmvae_gemm_tc_set_mode(mode)
        out = ops.gemm(a_s, b_s, None, M, N, K, a_k, b_k)
        ms = bench(lambda: ops.gemm(a_s, b_s, None, M, N, K, a_k, b_k))
        err = ((out.double() - ref).norm() / ref.norm()).item() if ref is not None else float("nan")
        res[name] = (err, ms, 2.0 * M * N * K / ms / 1e9)
    lib.mmvae_gemm_tc_set_mode(0)
    print(f"M={M} N={N} K={K} a_k={a_k} b_k={b_k}: " + "  ".join(f"{k}: err={v[0]:.2e} {v[1]:.3f}ms {v[2]:.1f}TF" for k, v in res.items()), flush=True)
