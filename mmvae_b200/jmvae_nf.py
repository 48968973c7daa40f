"""JMVAE / JMVAE-NF: joint encoder q(z|x,y), shared decoders, unimodal posteriors q(z|x_m) given by a Gaussian pushed
through a MAF / IAF flow and fitted to the joint posterior.

Reference: models/jmvae_nf/jmvae_nf.py:20-162 (class, forward, compute_kld, compute_joint_ll_from_uni,
compute_recon_loss), :189-193 (step), :209-270 (compute_joint_likelihood); objectives.py:177-220 (m_jmvae_nf);
models/jmvae_nf/jmvae_nf_mnist_svhn_dcca.py:33-94 (MNIST-SVHN subclass).

Every contraction, sampling, density and reduction below is a libmmvae_b200 kernel (ops.*); torch is used for
autograd bookkeeping, views and the final scalar sums.
"""
import math

import numpy as np
import torch
import torch.distributions as dist
import torch.nn as nn

from . import ops
from .models import Multi_VAES, dist_dict
from .networks import (Decoder_AE_MLP, Decoder_VAE_SVHN, DoubleHeadJoint, Encoder_VAE_MLP, Encoder_VAE_SVHN,
                       LinearCCAEncoder, TwoStepsEncoder)
from .vaes import VAE_IAF_Config, VAE_MAF_Config, VAEConfig, draw_noise, my_VAE, my_VAE_IAF, my_VAE_MAF

LOG2PI = math.log(2 * math.pi)
NEG_LOSS = {"normal": "neg_mse", "laplace": "neg_l1", "bernoulli": "neg_bce"}       # recon_loss_dict, objectives.py:177


class JMVAE_NF(Multi_VAES):
    """jmvae_nf.py:20-39."""

    def __init__(self, params, joint_encoder, vaes):
        super().__init__(params, vaes)
        self.joint_encoder = joint_encoder
        self.qz_xy = dist_dict[params.dist]
        self.qz_xy_params = None                      # populated in forward
        self.beta_kl = params.beta_kl
        self.fix_jencoder = params.fix_jencoder
        self.fix_decoders = params.fix_decoders
        self.lik_scaling = (1, 1)
        self.decrease_beta_kl = params.decrease_beta_kl
        self.ratio_kl_recon = [None, None]
        self.no_recon = getattr(params, "no_recon", False)
        self.train_latents = None
        self.linear_warmup = getattr(params, "linear_warmup", False)

    # ---- sampling from the joint posterior (one fused rsample launch) ---------------------------------------
    def _joint_sample(self, x, noise=None, K=1, bk_layout=False):
        """-> mu [B,L], std [B,L], z [K,B,L] ([B,K,L] when bk_layout)."""
        mu, raw = self.joint_encoder.forward_raw(x)
        B, L = mu.shape
        shape = (1, B, K, L) if bk_layout else (1, K, B, L)
        if noise is None:
            noise = draw_noise("normal", shape, mu.device)
        z, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), noise.reshape(shape), "normal", "exp_half_eps",
                                   bk_layout)
        return mu, std.reshape(B, L), z[0]

    def forward(self, x, noise=None):
        """jmvae_nf.py:41-52 -> (qz_xy, recons, z_xy)."""
        mu, std, z = self._joint_sample(x, noise)
        z_xy = z[0]
        self.qz_xy_params = (mu, std)
        qz_xy = self.qz_xy(mu, std, validate_args=False)
        recons = [vae.decoder(z_xy)["reconstruction"] for vae in self.vaes]
        return qz_xy, recons, z_xy

    # ---- regulariser ----------------------------------------------------------------------------------------
    def compute_kld(self, x, noise=None, noise_uni=None):
        """jmvae_nf.py:56-85: sum_m [ beta_kl * (log q(z|x,y) - log q_m(z|x_m)) (+ recon_m) ] on a fresh z ~ q(z|x,y)
        (the reference's second stochastic forward; its discarded decodes are skipped — SURVEY A14)."""
        mu, std, z = self._joint_sample(x, noise)
        z_xy = z[0]
        self.qz_xy_params = (mu.detach(), std.detach())       # no autograd graph kept alive between steps (CUDA-graph capture)
        lq_xy = ops.GaussRows.apply(z_xy, mu, std, "std", 1).sum()
        reg, details = 0, {}
        for m, vae in enumerate(self.vaes):
            z0, ladj = vae.flow_forward(z_xy)
            mu_m, lv_m = vae.encode_raw(x[m])
            log_q_z0 = ops.GaussRows.apply(z0, mu_m, lv_m, "log_var", 1)
            tot = log_q_z0.sum() if ladj is None else (log_q_z0 + ladj).sum()
            details[f"kld_{m}"] = lq_xy - tot
            if self.no_recon:
                reg = reg + self.beta_kl * details[f"kld_{m}"]
            else:
                out = vae(x[m], noise=None if noise_uni is None else noise_uni[m])
                details[f"recon_loss_{m}"] = self.compute_recon_loss(x[m], out.recon_x, m)
                self.ratio_kl_recon[m] = 1
                reg = reg + (self.beta_kl * details[f"kld_{m}"] + self.ratio_kl_recon[m] * details[f"recon_loss_{m}"])
        return reg, details

    def _sq_err_sum(self, a, b):
        """F.mse_loss(a, b, reduction='sum') over [B, D] rows, through the log-likelihood reduction kernel."""
        B = a.shape[0]
        a2, b2 = a.reshape(B, -1), b.reshape(B, -1)
        D = a2.shape[1]
        if D % 4:                                     # the reduction kernel reads 128-bit words
            pad = 4 - D % 4
            a2 = torch.cat([a2, torch.zeros(B, pad, device=a.device)], 1)
            b2 = torch.cat([b2, torch.zeros(B, pad, device=a.device)], 1)
        return -ops.LogLik.apply(a2, b2, 1, "neg_mse", 1.0, False)[0].sum()

    def compute_recon_loss(self, x, recon, m):
        """jmvae_nf.py:147-162: squared error in data space, or in the (frozen) DCCA embedding space."""
        if hasattr(self, "dcca") and getattr(self.params, "dcca", False):
            self.dcca[m].requires_grad_(False)
            with torch.no_grad():
                t = self.dcca[m](x).embedding
            recon_t = self.dcca[m](recon).embedding
            return self._sq_err_sum(recon_t, t)
        return self._sq_err_sum(recon, x)

    def step(self, epoch):
        """jmvae_nf.py:189-193."""
        if epoch >= self.params.warmup:
            self.beta_kl *= self.decrease_beta_kl

    # ---- the objective (objectives.m_jmvae_nf calls this) ---------------------------------------------------------
    def objective(self, x, epoch=1, warmup=0, beta_prior=1, noise=None):
        """objectives.py:179-220.  noise: None or dict(eps1=[B,L], eps2=[B,L], eps_uni=[M x [B,L]])."""
        noise = noise or {}
        if epoch >= warmup:
            self.joint_encoder.requires_grad_(not self.fix_jencoder)
            for vae in self.vaes:
                vae.decoder.requires_grad_(not self.fix_decoders)
        beta_reg = ((epoch - 1) / warmup if epoch - 1 <= warmup else 1) if self.linear_warmup else 1
        mu, std, z = self._joint_sample(x, noise.get("eps1"))
        z_xy = z[0]
        self.qz_xy_params = (mu.detach(), std.detach())       # no autograd graph kept alive between steps (CUDA-graph capture)
        loss, details = 0, {}
        for m, xm in enumerate(x):
            recon = self.vaes[m].decoder(z_xy)["reconstruction"]
            assert recon.shape == xm.shape, f"Sizes are different : {recon.shape, xm.shape}"
            B = xm.shape[0]
            neg = ops.LogLik.apply(recon.reshape(B, -1), self._flat_x(xm), 1, NEG_LOSS[self.params.recon_losses[m]],
                                   float(self.lik_scaling[m]), False)[0].sum()
            details[f"loss_{m}"] = -neg
            loss = loss + neg
        details["loss"] = loss
        details["kld_prior"] = ops.GaussRows.apply(None, mu, std, "kl_prior", 1).sum()
        if epoch >= warmup or self.linear_warmup:
            reg, det = self.compute_kld(x, noise.get("eps2"), noise.get("eps_uni"))
            details["reg"] = reg
            details.update(det)
        else:
            details["reg"] = 0
        return loss - beta_reg * (beta_prior * details["kld_prior"] + details["reg"]), details

    # ---- evaluators -------------------------------------------------------------------------------------------
    @torch.no_grad()
    def compute_joint_likelihood(self, data, K=1000, batch_size_K=100, noise=None):
        """jmvae_nf.py:209-270: z ~ q(z|x,y); weights p(x|z)p(y|z)p(z)/q(z|x,y).  noise: (K,B,L)."""
        mu, std, z = self._joint_sample(data, noise, K=K)
        self.qz_xy_params = (mu, std)
        z = z.permute(1, 0, 2)[:, :self._used(K, batch_size_K)].contiguous()
        ll = self._is_loglik(z, data, range(self.mod), "single", mu.unsqueeze(0), std.unsqueeze(0), K_total=K)
        self.last_per_datapoint = ll
        return {"likelihood": ll.mean()}

    @torch.no_grad()
    def compute_joint_ll_from_uni(self, data, cond_mod, K=1000, batch_size_K=100, noise=None):
        """jmvae_nf.py:87-143: z ~ q(z|x_cond) (through the flow); log q = log N(z0; mu, e^{lv}) - log|det J|.
        noise: (B,Kused,L)."""
        Ku = self._used(K, batch_size_K)
        z, z0, ladj, mu, raw, _ = self._cond_latents(cond_mod, data[cond_mod], Ku, noise)
        B, _, L = z.shape
        lq = ops.GaussRows.apply(z0.reshape(B * Ku, L), mu, raw, "log_var", Ku)
        if ladj is not None:
            lq = lq - ladj
        ll = self._is_loglik(z.contiguous(), data, range(self.mod), "external", lq_ext=lq.reshape(B, Ku), K_total=K)
        return {f"joint_ll_from_{cond_mod}": ll.mean()}


    # ---- product-of-experts subset posteriors (jmvae_nf.py:272-437) ----------------------------------------------------
    @torch.no_grad()
    def sample_from_moe_subset(self, subset, data, indices=None, noise=None):
        """jmvae_nf.py:272-291: one latent per datapoint from the mixture of the subset's unimodal posteriors.
        indices: the modality drawn for each datapoint (np.random.choice(subset, n) when None); noise: one (n_m, L)
        tensor per subset member."""
        n = len(data[0])
        if indices is None:
            indices = np.random.choice(subset, size=n)
        indices = torch.as_tensor(np.asarray(indices), device=data[0].device)
        zs = torch.zeros((n, self.params.latent_dim), device=data[0].device)
        for j, m in enumerate(subset):
            sel = indices == m
            if bool(sel.any()):
                zs[sel] = self.vaes[m].sample(data[m][sel].contiguous(), noise=None if noise is None else noise[j], decode=False).z
        return zs

    @torch.no_grad()
    def poe_experts(self, subset, data):
        """(mu_m, log_var_m) of the subset's encoders.  They do not depend on z: the sampler evaluates them ONCE, where the
        reference re-runs every encoder at each of the mcmc_steps * (n_lf + 1) density evaluations (jmvae_nf.py:317)."""
        return {m: self.vaes[m].encode_raw(data[m]) for m in subset}

    def compute_poe_posterior(self, subset, z_, data, divide_prior=True, grad=True, experts=None):
        """jmvae_nf.py:294-329: ln q(z|X_s) = sum_{m in s} [ln N(flow_m(z); mu_m, e^{lv_m}) + ln|det J_m|] - ln p(z) (when
        divide_prior) -> (values (n,), d values.sum()/dz) when grad, else values."""
        z = z_.clone().detach().requires_grad_(grad)
        n, L = z.shape
        with torch.set_grad_enabled(grad or torch.is_grad_enabled()):
            lnq = 0
            if divide_prior:
                lnq = lnq + (0.5 * (z * z + LOG2PI)).sum(dim=1)
            for m in subset:
                vae = self.vaes[m]
                z0, ladj = vae.flow_forward(z)
                mu, lv = experts[m] if experts is not None else vae.encode_raw(data[m])
                lq = ops.GaussRows.apply(z0, mu, lv, "log_var", 1)
                lnq = lnq + (lq if ladj is None else lq + ladj)
            if grad:
                g = torch.autograd.grad(lnq.sum(), z)[0]
                return lnq.detach(), g
        return lnq

    def sample_from_poe_subset(self, subset, data, ax=None, mcmc_steps=100, n_lf=10, eps_lf=0.01, K=1, divide_prior=True,
                               indices=None, noise=None, gammas=None, accepts=None, graph=None):
        """jmvae_nf.py:332-437: Hamiltonian Monte-Carlo on ln q(z|X_s), K chains per datapoint started from the mixture of
        experts -> z (K, n_data, L).  `ax` (trajectory plot) is ignored.  indices / noise feed sample_from_moe_subset;
        gammas [steps x (K*n,L)] are the momenta and accepts [steps x (K*n,)] the uniforms of the accept test (drawn on
        the device when None).  graph (default: on when every draw is internal or supplied as tensors): one leapfrog step
        — density, its z-gradient through the flows, the two half kicks and the drift — is captured in a CUDA graph and
        replayed mcmc_steps * n_lf times, as the ~100 launches of a step are otherwise launch-latency-bound."""
        n_data = len(data[0])
        data = [torch.cat([d] * K) for d in data]
        n = len(data[0])
        dev = data[0].device
        L = self.params.latent_dim
        z0 = self.sample_from_moe_subset(subset, data, indices, noise)
        experts = self.poe_experts(subset, data)
        acc_nbr = torch.zeros(n, 1, device=dev)
        graph = (dev.type == "cuda") if graph is None else graph

        def density(zz):
            return self.compute_poe_posterior(subset, zz, data, divide_prior=divide_prior, grad=True, experts=experts)

        zb, rb = torch.empty(n, L, device=dev), torch.empty(n, L, device=dev)       # static buffers of the leapfrog step
        gb, lb = torch.empty(n, L, device=dev), torch.empty(n, device=dev)
        half = eps_lf / 2

        def leapfrog():
            rho_ = rb + half * gb                      # step 1 (rho - eps/2 * (-g))
            zb.add_(rho_, alpha=eps_lf)                # step 2
            lnq, g = density(zb)
            gb.copy_(g)
            lb.copy_(lnq)
            rb.copy_(rho_ + half * g)                  # step 3

        cg = None
        if graph and n_lf > 0:
            zb.copy_(z0)
            rb.zero_()
            lb_, g_ = density(zb)
            gb.copy_(g_)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):                     # warm the allocator / autograd on the side stream
                    leapfrog()
            torch.cuda.current_stream().wait_stream(side)
            cg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cg):
                leapfrog()

        z = z0
        for i in range(mcmc_steps):
            rho = torch.randn_like(z) if gammas is None else gammas[i].to(dev)
            lnq, g = density(z)
            H0 = -lnq + 0.5 * torch.norm(rho, dim=1) ** 2
            zb.copy_(z)
            rb.copy_(rho)
            gb.copy_(g)
            lb.copy_(lnq)
            for _ in range(n_lf):
                if cg is not None:
                    cg.replay()
                else:
                    leapfrog()
            H = -lb + 0.5 * torch.norm(rb, dim=1) ** 2
            alpha = torch.exp(H0 - H)
            u = torch.rand(n, device=dev) if accepts is None else accepts[i].to(dev)
            moves = (u < alpha).to(torch.float32).reshape(n, 1)
            acc_nbr += moves
            z = zb * moves + (1 - moves) * z0
            z0 = z
        self.last_acceptance = acc_nbr / max(mcmc_steps, 1)
        return z.detach().reshape(K, n_data, L)


class JMVAE_NF_DCCA_MNIST_SVHN(JMVAE_NF):
    """jmvae_nf_mnist_svhn_dcca.py:33-94.  The DCCA first-stage weights are not shipped with the reference
    (`load_dcca_mnist_svhn` reads ../experiments/dcca/...): pass them as `dcca=[enc0, enc1]` (LinearCCAEncoder) or get
    randomly initialised, frozen ones of the same shape."""
    shape_mods = [(1, 28, 28), (3, 32, 32)]
    mod = 2
    modelName = "jmvae_nf_dcca_mnist_svhn"

    def __init__(self, params, dcca=None):
        if params.no_nf:
            vae_config, vae = VAEConfig, my_VAE
        else:
            vae_config = VAE_IAF_Config if params.flow == "iaf" else VAE_MAF_Config
            vae = my_VAE_IAF if params.flow == "iaf" else my_VAE_MAF
        pre = [VAEConfig((1, 28, 28), 20), VAEConfig((3, 32, 32), 20)]
        joint_encoder = DoubleHeadJoint(512, pre[0], pre[1], Encoder_VAE_MLP, Encoder_VAE_SVHN, params)
        c1, c2 = vae_config((1, 28, 28), params.latent_dim), vae_config((3, 32, 32), params.latent_dim)
        if hasattr(params, "n_made_blocks"):
            c1.n_made_blocks = c2.n_made_blocks = params.n_made_blocks
        dcca_mods = None
        if getattr(params, "dcca", False):
            if dcca is None:
                f1, f2 = VAEConfig((1, 28, 28), 16), VAEConfig((3, 32, 32), 16)
                dcca = [LinearCCAEncoder(Encoder_VAE_MLP(f1), 16, params.dim_dcca),
                        LinearCCAEncoder(Encoder_VAE_SVHN(f2), 16, params.dim_dcca)]
                for d in dcca:
                    nn.init.eye_(d.w)
            dcca_mods = dcca
            enc1, enc2 = TwoStepsEncoder(dcca[0], params), TwoStepsEncoder(dcca[1], params)
        else:
            enc1, enc2 = Encoder_VAE_MLP(c1), Encoder_VAE_SVHN(c2)
        vaes = nn.ModuleList([vae(model_config=c1, encoder=enc1, decoder=Decoder_AE_MLP(c1)),
                              vae(model_config=c2, encoder=enc2, decoder=Decoder_VAE_SVHN(c2))])
        super().__init__(params, joint_encoder, vaes)
        if dcca_mods is not None:
            self.dcca = dcca_mods             # plain list, as the reference: the modules are owned by the encoders
            for d in dcca_mods:
                d.requires_grad_(False)
        self.vaes[0].modelName, self.vaes[1].modelName = "mnist", "svhn"
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = ((3 * 32 * 32) / (1 * 28 * 28), 1) if s == 0.0 else (s, 1)


VAE_jnf_mnist_svhn_dcca = JMVAE_NF_DCCA_MNIST_SVHN
