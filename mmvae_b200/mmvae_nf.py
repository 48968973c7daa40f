"""MMVAE-NF: mixture of experts whose unimodal posteriors are Gaussians pushed through an IAF — SURVEY.md §8(f) rank 4.

Reference: models/mmvae_nf/mmvae_nf.py:20-88 (class, forward), models/mmvae_nf/mnist_svhn.py:35-68 (MNIST-SVHN
subclass: `my_VAE_IAF` on the pythae MLP / SVHN conv nets); objective `m_elbo_nf` (objectives.py:463-479).

What runs where: encoders / decoders through the GEMM and conv kernels (networks.py); z0 = mu + eps * exp(lv / 2) is the
rsample kernel; the IAF sampling direction is one MADE pass per block (flows.IAF.inverse); log q(z0) is the Gaussian
row-sum kernel; each decoder runs ONCE on the latents of every expert (M*B rows, `_run_branches` puts the modalities on
side streams) and the squared-error terms are the `neg_mse` mode of the log-likelihood reduction kernel.
"""
import math

import torch
import torch.nn as nn

from . import ops
from .models import Multi_VAES
from .networks import Decoder_AE_MLP, Decoder_VAE_SVHN, Encoder_VAE_MLP, Encoder_VAE_SVHN
from .vaes import VAE_IAF_Config, my_VAE_IAF

LOG2PI = math.log(2 * math.pi)


class MMVAE_NF(Multi_VAES):
    """mmvae_nf.py:20-88."""

    def __init__(self, params, vaes):
        super().__init__(params, vaes)
        assert params.dist == "normal"                 # mmvae_nf.py:26: Gaussian prior and posteriors

    def _latents(self, x, noise=None):
        """Per expert m: (ln q(z_m|x_m) [B], z_m [B,L]) with ln q = sum_L -1/2 (lv + (z0-mu)^2 e^{-lv}) - log|det J|
        (mmvae_nf.py:47-50: the Gaussian density WITHOUT its 2 pi term)."""
        def one(m, vae):
            o = vae.sample(x[m], noise=None if noise is None else noise[m], decode=False)
            L = o.mu.shape[1]
            lq = ops.GaussRows.apply(o.z0, o.mu, o.log_var, "log_var", 1) + 0.5 * L * LOG2PI - o.log_abs_det_jac
            return lq, o.z
        outs = self._run_branches([(lambda m=m, vae=vae: one(m, vae)) for m, vae in enumerate(self.vaes)],
                                  shared=tuple(x) + (() if noise is None else tuple(noise)))
        return [o[0] for o in outs], [o[1] for o in outs]

    def _decode_all(self, zs):
        """Decoder d on the latents of every expert at once -> recon_d [M*B, C,H,W] (row e*B+b = decoder_d(z_e[b]))."""
        rows = torch.cat(zs, dim=0)
        outs = self._run_branches([(lambda vae=vae: (vae.decoder(rows)["reconstruction"],)) for vae in self.vaes],
                                  shared=(rows,))
        return [o[0] for o in outs]

    def forward(self, x, noise=None):
        """mmvae_nf.py:30-62 -> (ln_qz_xs [M x (B,)], zs [M x (B,L)], recons[e][d]).  noise: M tensors (B,L)."""
        ln_q, zs = self._latents(x, noise)
        B = zs[0].shape[0]
        rec = self._decode_all(zs)
        recons = [[rec[d][e * B:(e + 1) * B] for d in range(self.mod)] for e in range(self.mod)]
        return ln_q, zs, recons

    def objective(self, x, noise=None):
        """objectives.py:463-479 (`m_elbo_nf`).  `row[e].sum()` at :471 indexes the (B,) tensor ln q(z_e|x_e) with the
        expert index e, so only DATAPOINT e's log-density enters the KL term; reproduced as is (needs B >= M)."""
        M = self.mod
        ln_q, zs = self._latents(x, noise)
        B = zs[0].shape[0]
        rows = torch.cat(zs, dim=0)

        def chain(d):
            recon = self.vaes[d].decoder(rows)["reconstruction"]
            lp, _ = ops.LogLik.apply(recon.reshape(M * B, -1), self._flat_x(x[d]), 1, "neg_mse",
                                     0.5 * self.lik_scaling[d] / M, False)
            return (lp,)

        elbo = sum(o[0].sum() for o in self._run_branches([(lambda d=d: chain(d)) for d in range(M)],
                                                          shared=(rows,) + tuple(x)))
        for e in range(M):
            kld = ln_q[e][e].sum() + 0.5 * torch.sum(zs[e] ** 2)
            elbo = elbo - kld / M
        return elbo

    def reconstruct(self, data, runPath=None, epoch=None):
        pass                                           # mmvae_nf.py:78-81

    def compute_joint_ll_from_uni(self, data, cond_mod, K=1000, batch_size_K=100):
        pass                                           # mmvae_nf.py:84-85: not defined for this model

    def compute_joint_likelihood(self, data, cond_mod=None, K=1000, batch_size_K=100):
        pass                                           # mmvae_nf.py:87-88


class MMVAE_NF_MNIST_SVHN(MMVAE_NF):
    """models/mmvae_nf/mnist_svhn.py:35-68."""
    shape_mod1, shape_mod2 = (1, 28, 28), (3, 32, 32)

    def __init__(self, params):
        c1, c2 = VAE_IAF_Config((1, 28, 28), params.latent_dim), VAE_IAF_Config((3, 32, 32), params.latent_dim)
        vaes = nn.ModuleList([my_VAE_IAF(c1, Encoder_VAE_MLP(c1), Decoder_AE_MLP(c1)),
                              my_VAE_IAF(c2, Encoder_VAE_SVHN(c2), Decoder_VAE_SVHN(c2))])
        super().__init__(params, vaes)
        self.modelName = "mmvae_nf_mnist_svhn"
        self.vaes[0].modelName, self.vaes[1].modelName = "mnist", "svhn"
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = ((3 * 32 * 32) / (1 * 28 * 28), 1) if s == 0.0 else (s, 1)

    def step(self, epoch):
        return


VAE_mmvae_nf_mnist_svhn = MMVAE_NF_MNIST_SVHN
