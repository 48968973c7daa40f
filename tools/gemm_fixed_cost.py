"""Ablation of the fixed cost of one tcgen05 GEMM launch (old dispatch gate, bit 32768) on small shapes: debug bits switch
off the epilogue stores (1), the operand loads (2, 4), the hi/lo split (8), all but one MMA (16); mode 1 = a single TF32
pass; bias / activation off; BN=32 shape for comparison."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops
from tools.gemm_tiny import bench  # noqa: E402
lib = _lib.load()
for M, N, K in ((128, 512, 64), (128, 128, 64), (128, 32, 64), (2048, 2048, 64)):
    A = torch.randn(M, K, device="cuda")
    B = torch.randn(N, K, device="cuda")
    bias = torch.randn(N, device="cuda")
    out = torch.empty(M, N, device="cuda")
    row = []
    for tag, mode, b, act in (("full", 0, bias, "relu"), ("no bias/act", 0, None, None), ("no stores", 1 << 8, bias, "relu"),
                              ("no global store instr", 64 << 8, bias, "relu"), ("no split", 8 << 8, bias, "relu"), ("one MMA", 16 << 8, bias, "relu"), ("1xTF32", 1, bias, "relu"),
                              ("no stores+no split+one MMA", (1 | 8 | 16) << 8, bias, "relu")):
        lib.mmvae_gemm_tc_set_mode(mode | (32768 << 8))
        row.append(f"{tag}: {bench(lambda: ops.gemm(A, B, b, M, N, K, 1, 1, act=act, out=out), n=20):.1f}")
    lib.mmvae_gemm_tc_set_mode(0)
    print(f"M={M} N={N} K={K}  " + " | ".join(row), flush=True)
