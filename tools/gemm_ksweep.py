"""GPU time of the tcgen05 GEMM against K (old dispatch gate, debug bit 32768) for a few (M, N): looks for a fixed cost."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops
from tools.gemm_tiny import bench  # noqa: E402  (runs its table first)
lib = _lib.load()
lib.mmvae_gemm_tc_set_mode(32768 << 8)
for M, N, ak, bk in ((128, 512, 1, 1), (512, 512, 0, 0), (2048, 2048, 1, 1), (7680, 512, 1, 1)):
    row = []
    for K in (32, 64, 96, 128, 160, 192, 256, 512, 1024, 2048):
        A = torch.randn((M, K) if ak else (K, M), device="cuda")
        B = torch.randn((N, K) if bk else (K, N), device="cuda")
        bias = torch.randn(N, device="cuda")
        out = torch.empty(M, N, device="cuda")
        row.append(f"K={K}: {bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act='relu', out=out), n=20):.1f}")
    print(f"M={M} N={N} a_k={ak} b_k={bk}  " + "  ".join(row), flush=True)
lib.mmvae_gemm_tc_set_mode(0)
