"""A/B of the tcgen05 GEMM's bf16x3 mode against 3xTF32 (debug bit 65536) on K-major x K-major shapes: time and error."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mmvae_b200 import ops, _lib
lib = _lib.load()
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for (M, N, K) in ((7680, 784, 512), (512, 20, 7680), (8192, 64, 512), (2048, 128, 1024), (7680, 512, 512), (122880, 128, 512)):
    A = torch.randn(M, K, device="cuda") * K ** -0.5; B = torch.randn(N, K, device="cuda"); b = torch.randn(N, device="cuda")
    ref = A.double() @ B.double().t() + b.double()
    for bits, name in ((0, "bf16x3"), (65536, "3xTF32")):
        lib.mmvae_gemm_tc_set_mode(bits << 8)
        n0 = lib.mmvae_gemm_tc_launch_count()
        out = ops.gemm(A, B, b, M, N, K, 1, 1, None)
        on_tc = lib.mmvae_gemm_tc_launch_count() != n0
        err = ((out.double() - ref).norm() / ref.norm()).item()
        us = t(lambda: ops.gemm(A, B, b, M, N, K, 1, 1, None))
        print(f"M={M} N={N} K={K} {name}: {us:7.1f} us  err {err:.2e}  tcgen05={on_tc}", flush=True)
    lib.mmvae_gemm_tc_set_mode(0)
