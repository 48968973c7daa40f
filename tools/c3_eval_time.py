"""K=1000 evaluators of a JMVAE-NF (MAF) MNIST-SVHN model (BASELINE config C4 on a C3 model): joint likelihood from each
unimodal posterior and both conditional likelihoods, with the fused sequential flow kernel and with the L-pass GEMM
formulation of the same flow.    python tools/c3_eval_time.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mmvae_b200 import flows                  # noqa: E402
from oracle import bivae_oracle as O          # noqa: E402
from oracle import synth                      # noqa: E402
from tests.helpers import build_jmvae_nf      # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
spec = O.spec_jnf_mnist_svhn(20, "maf", dcca=False)
sd = synth.make_state_dict(spec, 1)
model = build_jmvae_nf(spec, sd, no_recon=True, flow="maf")
x = [t.cuda() for t in synth.make_inputs(spec, B, seed=2)]


def run():
    out = {}
    for c in range(2):
        out.update(model.compute_joint_ll_from_uni(x, c, K=1000, batch_size_K=100))
    out.update(model.compute_conditional_likelihood(x, 0, 1, K=1000, batch_size_K=100)[0])
    out.update(model.compute_conditional_likelihood(x, 1, 0, K=1000, batch_size_K=100)[0])
    return out


res = {}
for fused in (True, False):
    flows.FUSED_SEQUENTIAL = fused
    torch.manual_seed(5)
    run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    torch.manual_seed(5)
    e0.record()
    res[fused] = run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"fused={fused}: {ms:.1f} ms for {B} datapoints x 4 evaluators (K=1000) = {B / ms * 1e3:.0f} datapoints/s")
flows.FUSED_SEQUENTIAL = True
for k in res[True]:
    print(f"  {k}: fused {float(res[True][k]):.4f}  gemm {float(res[False][k]):.4f}")
