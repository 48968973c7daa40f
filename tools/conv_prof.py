"""Runs only the SVHN decoder (the conv stack that dominates the C2 step) forward + backward on 7680 latent rows, so
that one ncu pass captures exactly its kernels.   python tools/conv_prof.py [passes]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import mmvae_b200  # noqa: F401
from mmvae_b200.networks import Decoder_VAE_SVHN
from mmvae_b200.vaes import VAEConfig

passes = int(sys.argv[1]) if len(sys.argv) > 1 else 2
torch.manual_seed(0)
dec = Decoder_VAE_SVHN(VAEConfig((3, 32, 32), 20)).cuda()
z = torch.randn(7680, 20, device="cuda", requires_grad=True)
g = torch.randn(7680, 3, 32, 32, device="cuda") * 1e-3
for i in range(passes):
    out = dec(z)["reconstruction"]
    out.backward(g)
    torch.cuda.synchronize()
print("ok", float(out.mean()))
