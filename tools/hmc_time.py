"""Times JMVAE_NF.sample_from_poe_subset (Hamiltonian sampler on the PoE posterior, jmvae_nf.py:332-437) on the GPU with
and without the CUDA-graph leapfrog, and the oracle port of the reference loop on the host cores (bounded sample).
    python tools/hmc_time.py [n_data] [K] [steps]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bivae_oracle as O          # noqa: E402
from oracle import synth                      # noqa: E402
from tests.helpers import build_jmvae_nf      # noqa: E402

n_data = int(sys.argv[1]) if len(sys.argv) > 1 else 100
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
spec = O.spec_jnf_mnist_svhn(20, "maf", dcca=False)
sd = synth.make_state_dict(spec, 1)
model = build_jmvae_nf(spec, sd, no_recon=True, flow="maf")
x = synth.make_inputs(spec, n_data, seed=2)
xc = [t.cuda() for t in x]
for graph in (False, True):
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        z = model.sample_from_poe_subset([0, 1], xc, mcmc_steps=steps, n_lf=10, eps_lf=0.01, K=K, graph=graph)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    evals = steps * 11
    print(f"gpu graph={graph}: {dt * 1e3:.1f} ms for {n_data * K} chains x {steps} steps ({dt / evals * 1e6:.0f} us per density+gradient "
          f"evaluation), acceptance {float(model.last_acceptance.mean()):.3f}")
cs = max(1, steps // 20)
n = n_data * K
gen = torch.Generator().manual_seed(3)
gam = [torch.randn(n, 20, generator=gen) for _ in range(cs)]
acc = [torch.rand(n, generator=gen) for _ in range(cs)]
t0 = time.perf_counter()
O.jmvae_nf_hmc_poe_subset(spec, sd, [0, 1], x, K, torch.randn(n, 20, generator=gen), gam, acc, 10, 0.01, True)
dt = time.perf_counter() - t0
print(f"cpu port ({torch.get_num_threads()} threads): {dt * 1e3:.1f} ms for {cs} steps -> {dt / cs * steps * 1e3:.0f} ms per {steps} steps")
