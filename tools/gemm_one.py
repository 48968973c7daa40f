"""One tcgen05 GEMM shape launched a few times (for ncu --set full --import-source on).  python tools/gemm_one.py M N K"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib, ops
M, N, K = (int(v) for v in sys.argv[1:4])
_lib.load().mmvae_gemm_tc_set_mode(32768 << 8)
A, B, bias = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda"), torch.randn(N, device="cuda")
out = torch.empty(M, N, device="cuda")
for _ in range(6):
    ops.gemm(A, B, bias, M, N, K, 1, 1, act="relu", out=out)
torch.cuda.synchronize()
