"""Times the sequential direction of a 2-block MAF (sampling z = flow.inverse(z0), maf_model.py:78-110) on R rows: the fused
kernel (csrc/made_seq.cu) against the L-pass GEMM formulation.    python tools/flow_time.py [R] [L]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mmvae_b200 import flows  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
L = int(sys.argv[2]) if len(sys.argv) > 2 else 20
torch.manual_seed(0)
f = flows.MAF(L, 2, 3, 128).cuda()
x = torch.randn(R, L, device="cuda")


def timed(n=10):
    with torch.no_grad():
        for _ in range(3):
            f.inverse(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            f.inverse(x)
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for R_ in (R, 8 * R):
    x = torch.randn(R_, L, device="cuda")
    flows.FUSED_SEQUENTIAL = True
    a = timed()
    flows.FUSED_SEQUENTIAL = False
    b = timed(3)
    flows.FUSED_SEQUENTIAL = True
    mac = 2 * (L * 128 + 2 * 128 * 128 + 2 * L * 128)
    print(f"R={R_} L={L}: fused {a * 1e3:.0f} us ({R_ * mac * 2 / a / 1e9:.2f} TFLOP/s of one-MADE-pass-per-block work), "
          f"GEMM formulation {b * 1e3:.0f} us ({b / a:.1f}x)")
