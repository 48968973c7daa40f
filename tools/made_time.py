"""GPU time of one MADE (20 -> 128 x3 -> 40, the flow block of config C3) forward and forward+backward on R rows, fused kernels
(csrc/made_train.cu) against the per-layer formulation; CUDA-graph replay of 20 back-to-back passes.  python tools/made_time.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                   # noqa: E402
from mmvae_b200 import flows, ops              # noqa: E402
from tools.gemm_probe import bench             # noqa: E402

for R in (128, 512, 1024):
    torch.manual_seed(0)
    made = flows.MADE(20, [128] * 3).cuda()
    x = torch.randn(R, 20, device="cuda", requires_grad=True)
    gy = torch.randn(R, 40, device="cuda")
    params = list(made.parameters())

    def fwd():
        with torch.no_grad():
            return made.raw(x)

    def fwd_bwd():
        y = made.raw(x)
        return torch.autograd.grad(y, [x] + params, gy)
    row = []
    for fused in (True, False):
        ops.MADE_FUSED = fused
        row.append(f"{'fused' if fused else 'per-layer'}: fwd {bench(fwd, n=20, graph=True):6.1f} us, fwd+bwd {bench(fwd_bwd, n=20, graph=True):6.1f} us")
    ops.MADE_FUSED = True
    print(f"R={R:5d}  " + "   ".join(row), flush=True)
