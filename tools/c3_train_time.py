"""Training step of BASELINE config C3 (JMVAE-NF, MAF flows, DCCA two-step encoders, m_jmvae_nf, B=128, K=1,
mnist_svhn/jmvae_nf_dcca.json) through train.Trainer: eager launches vs CUDA-graph replay.    python tools/c3_train_time.py"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mmvae_b200                             # noqa: E402
from mmvae_b200 import ops                    # noqa: E402
from mmvae_b200.train import Trainer          # noqa: E402

if "--simt" in sys.argv:
    ops.ALLOW_TC = False                       # every GEMM on the exact-fp32 SIMT kernel
for fixed in (True, False):
    params = argparse.Namespace(latent_dim=20, dist="normal", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                num_hidden_layers=1, beta_kl=1.0, decrease_beta_kl=1.0, fix_jencoder=fixed, fix_decoders=fixed,
                                no_recon=True, warmup=0, dcca=True, dim_dcca=9, flow="maf", no_nf=False, device="cuda",
                                data_path="", epochs=1)
    torch.manual_seed(1)
    model = mmvae_b200.VAE_jnf_mnist_svhn_dcca(params).cuda()
    tr = Trainer(model, "m_jmvae_nf", K=1, lr=1e-4, distributed=False)
    g = torch.Generator().manual_seed(3)
    x = [torch.rand(128, 1, 28, 28, generator=g).cuda(), torch.rand(128, 3, 32, 32, generator=g).cuda()]

    def timed(fn, n):
        for _ in range(5):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    n0 = ops.N_CALLS
    tr.step(x)
    calls = ops.N_CALLS - n0
    if "--detail" in sys.argv:
        model.branch_streams = False
        ws, tr._wgrad_stream = tr._wgrad_stream, None
        tr.step(x)
        torch.cuda.synchronize()
        ops.PROFILE = []
        tr.step(x)
        torch.cuda.synchronize()
        rec, ops.PROFILE = ops.PROFILE, None
        model.branch_streams, tr._wgrad_stream = True, ws
        agg = {}
        for name, work, e0, e1 in rec:
            a = agg.setdefault(name, [0.0, 0])
            a[0] += e0.elapsed_time(e1)
            a[1] += 1
        tot = sum(a[0] for a in agg.values())
        print(f"  kernel time {tot:.3f} ms over {len(rec)} calls")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:30]:
            print(f"  {k:30s} {a[0] * 1e3:8.1f} us {a[1]:4d} calls")
    eager = timed(lambda: tr.step(x), 30)
    tr.capture(x)
    graph = timed(lambda: tr.step_graph(x), 200)
    what = "second training phase (joint encoder and decoders frozen: unimodal encoders + flows train)" if fixed else \
        "everything trainable"
    print(f"C3 step, {what}: {calls} C-ABI calls; eager {eager:.3f} ms, graph replay {graph:.3f} ms "
          f"({128 / graph * 1e3:.0f} samples/s), loss {float(tr._static_loss):.3f}")
