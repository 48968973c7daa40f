"""MVAE (product of experts with the prior) on the same kernels as the MMVAE path — SURVEY.md §8(f) rank 4.

Reference: models/mvae/mvae.py:13-266 (`poe`, `forward`, `compute_joint_ll_from_uni`, `compute_joint_likelihood`),
models/mvae/mnist_svhn.py:22-45, models/mvae/msf.py:28-57; objective `m_self_built` (objectives.py:481-483).

What runs where: encoders / decoders through the GEMM and conv kernels (networks.py); the M unimodal PoE draws and the
joint draw are ONE rsample launch ((M+1) x B latents; std = exp(lnV/2) is the kernel's `exp_half` transform of the PoE
log-variance); each decoder runs ONCE on its unimodal and the joint latents (2B rows); the squared-error
reconstruction terms are the `neg_mse` mode of the log-likelihood reduction kernel; the K-sample evaluators reuse the
importance-weight / logsumexp kernels of the MMVAE evaluators (`qmode='single'`).  The PoE combination and the closed-form
KL act on (M+1) x B x L elements and stay torch glue.
"""
import math

import torch
import torch.distributions as dist
import torch.nn as nn

from . import ops
from .models import Multi_VAES
from .networks import Decoder_AE_MLP, Decoder_VAE_SVHN, Encoder_VAE_MLP, Encoder_VAE_SVHN
from .vaes import VAEConfig, draw_noise, my_VAE


class MVAE(Multi_VAES):
    """models/mvae/mvae.py:13-266."""

    def __init__(self, params, vaes):
        assert params.dist == "normal"                        # mvae.py:16: Gaussian prior and posteriors
        super().__init__(params, vaes)
        self.qz_x = dist.Normal
        self.subsampling = False
        self.subsets = []

    # ---- product of experts ------------------------------------------------------------------------------------
    @staticmethod
    def _poe_lnv(mus, log_vars):
        """mvae.py:27-45 -> (joint_mu, lnV) with joint_std = exp(lnV / 2); the prior N(0,1) is one more expert."""
        lnT = torch.stack([-l for l in log_vars] + [torch.zeros_like(log_vars[0])])
        lnV = -torch.logsumexp(lnT, dim=0)
        mu = (torch.exp(lnT[:-1]) * torch.stack(list(mus))).sum(dim=0) * torch.exp(lnV)
        return mu, lnV

    def poe(self, mus_list, log_vars_list, eps=1e-8):
        mu, lnV = self._poe_lnv(mus_list, log_vars_list)
        return mu, torch.exp(0.5 * lnV)

    def kl(self, mu, std):
        """kl_divergence(N(mu,std), N(*pz_params)).sum()  (mvae.py:60-61; torch _kl_normal_normal)."""
        ploc, pscale = self.pz_params[0], self.pz_params[1]
        vr = (std / pscale).pow(2)
        t1 = ((mu - ploc) / pscale).pow(2)
        return (0.5 * (vr + t1 - 1 - vr.log())).sum()

    def _encode(self, x):
        outs = self._run_branches([(lambda m=m, vae=vae: tuple(vae.encode_raw(x[m]))) for m, vae in enumerate(self.vaes)],
                                  shared=tuple(x))
        return [o[0] for o in outs], [o[1] for o in outs]

    def infer_latent_from_mod(self, cond_mod, x, noise=None):
        """mvae.py:64-71."""
        mu_u, lv_u = self.vaes[cond_mod].encode_raw(x)
        mu, lnV = self._poe_lnv([mu_u], [lv_u])
        B, L = mu.shape
        if noise is None:
            noise = draw_noise("normal", (B, L), mu.device)
        z, _ = ops.RSample.apply(mu.unsqueeze(0), lnV.unsqueeze(0), noise.reshape(1, 1, B, L), "normal", "exp_half", False)
        return z.reshape(B, L)

    # ---- forward = the training objective (mvae.py:74-143, subsampling off) -----------------------------------------
    def forward(self, x, noise=None):
        """-> dict(elbo, z_joint, joint_mu, joint_std).  noise: M+1 tensors (B,L): the unimodal PoE draws, then the joint."""
        M = self.mod
        mus, lvs = self._encode(x)
        B, L = mus[0].shape
        uni = [self._poe_lnv([mus[m]], [lvs[m]]) for m in range(M)]
        joint_mu, joint_lnV = self._poe_lnv(mus, lvs)
        mu_all = torch.stack([u[0] for u in uni] + [joint_mu])                 # [M+1,B,L]
        lnv_all = torch.stack([u[1] for u in uni] + [joint_lnV])
        if noise is None:
            noise = draw_noise("normal", (M + 1, 1, B, L), mu_all.device)
        elif not torch.is_tensor(noise):
            noise = torch.stack(list(noise))
        z, std = ops.RSample.apply(mu_all, lnv_all, noise.reshape(M + 1, 1, B, L), "normal", "exp_half", False)
        z = z.reshape(M + 1, B, L)

        def chain(d):                                                          # decoder d on [z_d ; z_joint]: 2B rows
            rows = torch.cat([z[d], z[M]], dim=0)
            recon = self.vaes[d].decoder(rows)["reconstruction"]
            lp, _ = ops.LogLik.apply(recon.reshape(2 * B, -1), self._flat_x(x[d]), 1, "neg_mse",
                                     0.5 * self.lik_scaling[d], False)       # -1/2 sum (x - recon)^2 * lik_scaling
            return (lp,)

        rec = sum(o[0].sum() for o in self._run_branches([(lambda d=d: chain(d)) for d in range(M)],
                                                         shared=(z,) + tuple(x)))
        elbo = rec - self.kl(mu_all, std)
        return dict(elbo=elbo, z_joint=z[M], joint_mu=joint_mu, joint_std=std[M])

    # ---- K-sample importance-sampled likelihoods -------------------------------------------------------------------
    @torch.no_grad()
    def compute_joint_likelihood(self, data, K=1000, batch_size_K=100, noise=None):
        """mvae.py:219-266: z ~ N(joint_mu, joint_std), K // batch_size_K chunks of batch_size_K draws, weights
        p(x,y|z) p(z) / q(z|x,y), ln K subtracted.  noise: (B, Kused, L)."""
        mus, lvs = self._encode(data)
        mu, lnV = self._poe_lnv(mus, lvs)
        B, L = mu.shape
        Ku = self._used(K, batch_size_K)
        if noise is None:
            noise = draw_noise("normal", (B, Ku, L), mu.device)
        z, std = ops.RSample.apply(mu.unsqueeze(0), lnV.unsqueeze(0), noise.reshape(1, B, Ku, L), "normal", "exp_half", True)
        ll = self._is_loglik(z.reshape(B, Ku, L), data, range(self.mod), "single", mu.unsqueeze(0), std, K_total=K)
        self.last_per_datapoint = ll
        return {"likelihood": ll.mean()}

    @torch.no_grad()
    def compute_joint_ll_from_uni(self, data, cond_mod, K=1000, batch_size_K=100, noise=None):
        """mvae.py:160-217: z ~ the ENCODER's Gaussian of the conditioning modality (the reference's own warning at
        :172-173: not the PoE with the prior).  noise: (K,B,L)."""
        vae = self.vaes[cond_mod]
        mu, raw = vae.encode_raw(data[cond_mod])
        B, L = mu.shape
        if noise is None:
            noise = draw_noise("normal", (K, B, L), mu.device)
        z, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), noise.reshape(1, K, B, L), "normal", "exp_half", False)
        z = z[0].permute(1, 0, 2)[:, :self._used(K, batch_size_K)].contiguous()
        ll = self._is_loglik(z, data, range(self.mod), "single", mu.unsqueeze(0), std, K_total=K)
        return {f"joint_ll_from_{cond_mod}": ll.mean()}


    # ---- product-of-experts subset posteriors (mvae.py:268-327) ---------------------------------------------------------
    @torch.no_grad()
    def sample_from_poe_subset(self, subset, data, K=1, divide_prior=True, noise=None):
        """mvae.py:268-301: K draws per datapoint from the PoE of the subset's experts and the prior -> (K, n, L).
        `poe` returns (mu, STD) and the reference samples with exp(0.5 * that) as the scale (:296-298); kept as is:
        the rsample kernel's `exp_half` transform is applied to the PoE standard deviation.  noise: (K, n, L)."""
        mus, lvs = [], []
        for m in subset:
            mu_u, lv_u = self.vaes[m].encode_raw(data[m])
            mus.append(mu_u)
            lvs.append(lv_u)
        joint_mu, joint_std = self.poe(mus, lvs)
        n, L = joint_mu.shape
        if noise is None:
            noise = draw_noise("normal", (K, n, L), joint_mu.device)
        z, _ = ops.RSample.apply(joint_mu.unsqueeze(0), joint_std.unsqueeze(0), noise.reshape(1, K, n, L), "normal",
                                 "exp_half", False)
        return z.reshape(K, n, L)

    def compute_poe_posterior(self, subset, z, data, divide_prior=True, grad=False, experts=None):
        """mvae.py:304-327: sum_{m in s} ln N(z; PoE(q_m, p)) - ln p(z) -> (n,).  The second value `poe` returns (a standard
        deviation) is read as a log-variance by the reference (:320-321); kept as is.  z may hold `div` rows per
        datapoint (row r belongs to datapoint r // div)."""
        R, L = z.shape
        lq = 0
        for m in subset:
            mu_u, lv_u = experts[m] if experts is not None else self.vaes[m].encode_raw(data[m])
            t0, t1 = self.poe([mu_u], [lv_u])
            lq = lq + ops.GaussRows.apply(z, t0, t1, "log_var", R // t0.shape[0])
        return lq + (0.5 * (z * z + math.log(2 * math.pi))).sum(dim=1)


class MVAE_MNIST_SVHN(MVAE):
    """models/mvae/mnist_svhn.py:22-45."""
    shape_mods = [(1, 28, 28), (3, 32, 32)]

    def __init__(self, params):
        c1, c2 = VAEConfig((1, 28, 28), params.latent_dim), VAEConfig((3, 32, 32), params.latent_dim)
        vaes = nn.ModuleList([my_VAE(c1, Encoder_VAE_MLP(c1), Decoder_AE_MLP(c1)),
                              my_VAE(c2, Encoder_VAE_SVHN(c2), Decoder_VAE_SVHN(c2))])
        super().__init__(params, vaes)
        self.modelName = "mvae_mnist_svhn"
        self.vaes[0].modelName, self.vaes[1].modelName = "mnist", "svhn"
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = ((3 * 32 * 32) / (1 * 28 * 28), 1.0) if s == 0 else (float(s), 1.0)


class MVAE_MNIST_SVHN_FASHION(MVAE):
    """models/mvae/msf.py:28-57 (subsampling of expert subsets off, as in the shipped config)."""
    shape_mods = [(1, 28, 28), (3, 32, 32), (1, 28, 28)]

    def __init__(self, params):
        c1, c2, c3 = (VAEConfig(s, params.latent_dim) for s in self.shape_mods)
        vaes = nn.ModuleList([my_VAE(c1, Encoder_VAE_MLP(c1), Decoder_AE_MLP(c1)),
                              my_VAE(c2, Encoder_VAE_SVHN(c2), Decoder_VAE_SVHN(c2)),
                              my_VAE(c3, Encoder_VAE_MLP(c3), Decoder_AE_MLP(c3))])
        super().__init__(params, vaes)
        self.modelName = "mvae_msf"
        for v, n in zip(self.vaes, ("mnist", "svhn", "fashion")):
            v.modelName = n
        r = (3 * 32 * 32) / (1 * 28 * 28)
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = (r, 1.0, r) if s == 0 else (float(s), 1.0, float(s))

    @torch.no_grad()
    def compute_conditional_likelihoods(self, data, K=1000, batch_size_K=100, noises=None):
        """msf.py:101-108: the pairwise conditionals plus `cond_poe_ll_{g}` from the two-modality PoE posteriors."""
        from .poe_subset import compute_all_cond_ll_from_poe_subsets
        d = super().compute_conditional_likelihoods(data, K, batch_size_K, noises)
        d.update(compute_all_cond_ll_from_poe_subsets(self, data, K, batch_size_K))
        return d


VAE_mvae_mnist_svhn = MVAE_MNIST_SVHN
VAE_mvae_msf = MVAE_MNIST_SVHN_FASHION


# ---------------------------------------------------------------------------------------------------------------
# MoPoE — models/moepoe/moepoe.py
# ---------------------------------------------------------------------------------------------------------------
class MOEPOE(Multi_VAES):
    """Mixture of products of experts (models/moepoe/moepoe.py:75-214).  The mixture components are the M unimodal
    posteriors and the PoE of every subset of >= 2 modalities (only the full subset includes the prior); one latent per
    datapoint is drawn from the component its batch position selects (mixture_component_selection, :20-40)."""

    def __init__(self, params, vaes):
        super().__init__(params, vaes)
        self.lik_scaling = tuple(1.0 for _ in vaes)               # moepoe.py:80 (subclasses set the pixel ratio)
        self.qz_x = dist.Normal

    @staticmethod
    def _components(mus, log_vars):
        """compute_poe_for_all_subsets, moepoe.py:44-70, appended to the unimodal experts (:93-100)."""
        from itertools import combinations
        M = len(mus)
        out_mu, out_lv = list(mus), list(log_vars)
        for k in range(2, M + 1):
            for tup in combinations(range(M), k):
                ms, ls = [mus[i] for i in tup], [log_vars[i] for i in tup]
                if k == M:
                    ms, ls = ms + [torch.zeros_like(mus[0])], ls + [torch.zeros_like(log_vars[0])]
                lnT = torch.stack([-l for l in ls])
                lnV = -torch.logsumexp(lnT, dim=0)
                out_mu.append((torch.exp(lnT) * torch.stack(ms)).sum(dim=0) * torch.exp(lnV))
                out_lv.append(lnV)
        return out_mu, out_lv

    @staticmethod
    def _select(mus, log_vars):
        C, B = len(mus), mus[0].shape[0]
        step = B // C
        bounds = [(k * step, B if k == C - 1 else (k + 1) * step) for k in range(C)]
        return (torch.cat([mus[k][a:b] for k, (a, b) in enumerate(bounds)]),
                torch.cat([log_vars[k][a:b] for k, (a, b) in enumerate(bounds)]))

    def forward(self, x, noise=None):
        """moepoe.py:83-135 -> dict(elbo, z_joint, mus, log_vars).  noise: (B,L)."""
        M = self.mod
        outs = self._run_branches([(lambda m=m, vae=vae: tuple(vae.encode_raw(x[m]))) for m, vae in enumerate(self.vaes)],
                                  shared=tuple(x))
        mus, lvs = self._components([o[0] for o in outs], [o[1] for o in outs])
        mu_r, lv_r = self._select(mus, lvs)
        B, L = mu_r.shape
        if noise is None:
            noise = draw_noise("normal", (B, L), mu_r.device)
        z, _ = ops.RSample.apply(mu_r.unsqueeze(0), lv_r.unsqueeze(0), noise.reshape(1, 1, B, L), "normal", "exp_half", False)
        z = z.reshape(B, L)

        def chain(d):
            recon = self.vaes[d].decoder(z)["reconstruction"]
            lp, _ = ops.LogLik.apply(recon.reshape(B, -1), self._flat_x(x[d]), 1, self.params.recon_losses[d],
                                     self.lik_scaling[d], False)
            return (lp,)

        elbo = sum(o[0].sum() for o in self._run_branches([(lambda d=d: chain(d)) for d in range(M)],
                                                          shared=(z,) + tuple(x)))
        mu_all, lv_all = torch.stack(mus), torch.stack(lvs)
        ploc, pscale = self.pz_params[0], self.pz_params[1]
        vr = torch.exp(lv_all) / pscale.pow(2)
        kld = 0.5 * (vr + ((mu_all - ploc) / pscale).pow(2) - 1 - vr.log())
        elbo = elbo - kld.sum() * getattr(self.params, "beta_kl", 1.0) / len(mus)
        return dict(elbo=elbo, z_joint=z, mus=mu_all, log_vars=lv_all)

    @torch.no_grad()
    def compute_joint_ll_from_uni(self, data, cond_mod, K=1000, batch_size_K=100, noise=None):
        """moepoe.py:160-214: as MVAE.compute_joint_ll_from_uni, but ln K is not subtracted (:212)."""
        vae = self.vaes[cond_mod]
        mu, raw = vae.encode_raw(data[cond_mod])
        B, L = mu.shape
        if noise is None:
            noise = draw_noise("normal", (K, B, L), mu.device)
        z, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), noise.reshape(1, K, B, L), "normal", "exp_half", False)
        z = z[0].permute(1, 0, 2)[:, :self._used(K, batch_size_K)].contiguous()
        ll = self._is_loglik(z, data, range(self.mod), "single", mu.unsqueeze(0), std, K_total=1)
        return {f"joint_ll_from_{cond_mod}": ll.mean()}

    def compute_joint_likelihood(self, data, K=1000, batch_size_K=100):
        raise NotImplementedError("the reference's MOEPOE.compute_joint_likelihood is unfinished (moepoe.py:217-245 "
                                  "raises ZeroDivisionError); there is nothing to be in parity with")


class MOEPOE_MNIST_SVHN(MOEPOE):
    """models/moepoe/mnist_svhn.py."""
    shape_mods = [(1, 28, 28), (3, 32, 32)]

    def __init__(self, params):
        c1, c2 = VAEConfig((1, 28, 28), params.latent_dim), VAEConfig((3, 32, 32), params.latent_dim)
        vaes = nn.ModuleList([my_VAE(c1, Encoder_VAE_MLP(c1), Decoder_AE_MLP(c1)),
                              my_VAE(c2, Encoder_VAE_SVHN(c2), Decoder_VAE_SVHN(c2))])
        super().__init__(params, vaes)
        self.modelName = "moepoe_mnist_svhn"
        self.vaes[0].modelName, self.vaes[1].modelName = "mnist", "svhn"
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = ((3 * 32 * 32) / (1 * 28 * 28), 1.0) if s == 0 else (float(s), 1.0)


VAE_moepoe_mnist_svhn = MOEPOE_MNIST_SVHN
