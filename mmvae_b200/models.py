"""Multimodal model classes with the reference API: Multi_VAES, MMVAE and the dataset
subclasses registered as VAE_<name>.  JMVAE_NF lives in jmvae_nf.py.

Reference: models/multi_vaes.py:32-59,219-355; models/mmvae/mmvae.py:24-234; models/mmvae/mmvae_mnist_svhn.py:33-54;
models/mmvae/mmvae_circles.py:33-59; models/mmvae/mnist_svhn_fashion.py:30-53.

`forward(x, K)` keeps the reference's 4-tuple.  The objectives and evaluators do not go through that tuple: they call
`_core(...)`, which encodes each modality ONCE (the reference tiles the batch K times, mmvae.py:41 — SURVEY A4),
draws all M*K*B latents with one rsample launch, runs each decoder once on all experts' latents, and reduces the
log-likelihoods / mixture weights in the fused kernels.
"""
import math

import torch
import torch.distributions as dist
import torch.nn as nn

from . import ops
from .networks import Decoder_AE_MLP, Decoder_VAE_SVHN, Encoder_VAE_MLP, Encoder_VAE_SVHN
from .vaes import VAEConfig, draw_noise, laplace_VAE, my_VAE

dist_dict = {"normal": dist.Normal, "laplace": dist.Laplace, "bernoulli": dist.Bernoulli}
EVAL_MAX_ROWS = 65536          # decoder rows per evaluator chunk (bounds the activation buffers to ~5 GB; 32768 measured 1.5 % slower at C4)


class Multi_VAES(nn.Module):
    """multi_vaes.py:32-59."""

    def __init__(self, params, vaes):
        super().__init__()
        self.mod = len(vaes)
        self.vaes = vaes
        self.modelName = None
        self.params = params
        self.data_path = getattr(params, "data_path", "")
        self._pz_params = nn.ParameterList([
            nn.Parameter(torch.zeros(1, params.latent_dim), requires_grad=False),
            nn.Parameter(torch.ones(1, params.latent_dim), requires_grad=False)])
        self.pz = dist_dict[params.dist]
        self.px_z = [dist_dict[r] for r in params.recon_losses]
        self.max_epochs = getattr(params, "epochs", 1)
        self.lik_scaling = tuple(1.0 for _ in vaes)
        self.eval_max_rows = EVAL_MAX_ROWS

    @property
    def pz_params(self):
        return self._pz_params

    # ---- independent per-modality chains on side streams -------------------------------------------------------
    # The M encoders, and the M decoders (+ their log-likelihood reductions), are independent chains of mostly small,
    # latency-bound launches.  With branch_streams=True (train.Trainer sets it) chain m < M-1 runs on its own side
    # stream and the last chain on the current stream, so the small launches of one modality hide under the large
    # kernels of another; autograd replays each backward node on the stream of its forward, so the backward overlaps the
    # same way.  Inside a CUDA-graph capture the fork/join events become graph edges.
    branch_streams = False

    def _run_branches(self, fns, shared=()):
        """fns: one closure per chain, each returning a tuple of tensors; shared: tensors the closures read that were
        produced on the current stream.  -> list of the closures' results, all usable on the current stream."""
        if not self.branch_streams or len(fns) < 2:
            return [f() for f in fns]
        cur = torch.cuda.current_stream()
        side = self.__dict__.setdefault("_side_streams", [])
        while len(side) < len(fns) - 1:
            side.append(torch.cuda.Stream())
        used = self.__dict__.setdefault("_side_used", set())
        outs = [None] * len(fns)
        for i, f in enumerate(fns[:-1]):
            s = side[i]
            s.wait_stream(cur)
            for t in shared:
                t.record_stream(s)
            with torch.cuda.stream(s):
                outs[i] = f()
            used.add(i)
        outs[-1] = fns[-1]()
        for i in range(len(fns) - 1):
            cur.wait_stream(side[i])
            for t in outs[i]:
                if torch.is_tensor(t):
                    t.record_stream(cur)
        return outs

    def join_branch_streams(self):
        """Makes the current stream wait for every side stream used since the last join (call after backward())."""
        used = self.__dict__.get("_side_used", set())
        for i in sorted(used):
            torch.cuda.current_stream().wait_stream(self._side_streams[i])
        used.clear()

    @property
    def latent_family(self):
        return self.params.dist

    # ---- shared evaluator pieces -------------------------------------------------------------------------
    @staticmethod
    def _used(K, batch_size_K):
        """`while stop_index <= K` drops the tail chunk; ln K is still subtracted."""
        return (K // batch_size_K) * batch_size_K

    def _flat_x(self, x_m):
        return ops._c(x_m.reshape(x_m.shape[0], -1), "data")

    def _lpx_rows(self, d, z_rows, x_d, x_div, scaled):
        """sum_pix log p(x_d | decoder_d(z)) for z_rows [R,L] -> [R]."""
        c = self.lik_scaling[d] if scaled else 1.0
        dec = self.vaes[d].decoder
        if not torch.is_grad_enabled() and hasattr(dec, "loglik_rows"):        # evaluators: the image never reaches HBM
            lp = dec.loglik_rows(z_rows, self._flat_x(x_d), x_div, self.params.recon_losses[d], c)
            if lp is not None:
                return lp
        recon = dec(z_rows)["reconstruction"]
        R = recon.shape[0]
        return ops.LogLik.apply(recon.reshape(R, -1), self._flat_x(x_d), x_div, self.params.recon_losses[d], c, False)[0]

    def _is_loglik(self, z, data, mods, qmode, mu=None, std=None, lq_ext=None, K_total=None, with_prior=True,
                   partial=False):
        """ln p_hat per datapoint from latents z [B,Ku,L]:  logsumexp_k( sum_{d in mods} log p(x_d|z_k)
        [+ log p(z_k) - q-term] ) - ln K_total, chunked over datapoints.  partial=True returns the per-datapoint
        (max, sum exp) pair of this K-shard instead (merged across ranks by parallel.combine_lse)."""
        B, Ku, L = z.shape
        out = torch.empty(B, device=z.device, dtype=torch.float32)
        out_se = torch.empty(B, device=z.device, dtype=torch.float32) if partial else None
        step = max(1, self.eval_max_rows // Ku)
        for b0 in range(0, B, step):
            b1 = min(B, b0 + step)
            zc = z[b0:b1].contiguous()
            rows = zc.reshape(-1, L)
            lpx = torch.stack([o[0] for o in self._run_branches(
                [(lambda d=d: (self._lpx_rows(d, rows, data[d][b0:b1], Ku, False).reshape(b1 - b0, Ku),)) for d in mods],
                shared=(rows,) + tuple(data[d] for d in mods))])
            if with_prior:
                a = ops.eval_logw(zc, None if mu is None else mu[:, b0:b1].contiguous(),
                                  None if std is None else std[:, b0:b1].contiguous(), lpx, self.pz_params[0],
                                  self.pz_params[1], self.params.dist, qmode,
                                  None if lq_ext is None else lq_ext[b0:b1].contiguous())
            else:
                a = lpx.sum(0) if len(mods) > 1 else lpx[0]
            mx, se = ops.lse_partial(a)
            if partial:
                out[b0:b1], out_se[b0:b1] = mx, se
            else:
                out[b0:b1] = ops.lse_finish(mx, se, K_total)
        return (out, out_se) if partial else out

    def _cond_latents(self, cond_mod, x_c, Ku, noise):
        """z ~ q(z|x_cond), Ku draws per datapoint -> [B,Ku,L] (flow-transformed for NF models)."""
        vae = self.vaes[cond_mod]
        mu, raw = vae.encode_raw(x_c)
        B, L = mu.shape
        if noise is None:
            noise = draw_noise(vae.dist, (B, Ku, L), mu.device)
        z0, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), noise.reshape(1, B, Ku, L), vae.dist, vae.xform,
                                    True)
        z0 = z0.reshape(B * Ku, L)
        z, ladj = vae.flow_inverse(z0)
        return z.reshape(B, Ku, L), z0.reshape(B, Ku, L), ladj, mu, raw, std.reshape(B, L)

    @torch.no_grad()
    def compute_conditional_likelihood(self, data, cond_mod, gen_mod, K=1000, batch_size_K=100, noise=None):
        """multi_vaes.py:271-318: ln p(x_gen|x_cond) ~ logsumexp_k log p(x_gen|z_k) - ln K, z_k ~ q(z|x_cond)."""
        Ku = self._used(K, batch_size_K)
        z = self._cond_latents(cond_mod, data[cond_mod], Ku, noise)[0]
        ll = self._is_loglik(z, data, [gen_mod], None, K_total=K, with_prior=False)
        return {f"cond_likelihood_{cond_mod}_{gen_mod}": ll.sum() / len(ll)}, ll.cpu()

    @torch.no_grad()
    def compute_conditional_likelihoods(self, data, K=1000, batch_size_K=100, noises=None):
        """multi_vaes.py:324-355 (incl. the M=3 mixture-of-two-sources subset conditional, :349-353)."""
        metrics, ll = {}, [[None] * self.mod for _ in range(self.mod)]
        for i in range(self.mod):
            for j in range(self.mod):
                if i != j:
                    nz = None if noises is None else noises[(j, i)]
                    m, per = self.compute_conditional_likelihood(data, j, i, K, batch_size_K, nz)
                    metrics.update(m)
                    ll[i][j] = per
        if self.mod == 3:
            for i in range(3):
                two = torch.stack([ll[i][j] for j in range(3) if j != i])
                metrics[f"cond_lw_subset_{i}"] = (torch.logsumexp(two, dim=0) - math.log(2)).mean()
        return metrics

    @torch.no_grad()
    def compute_uni_ll_from_prior(self, data, mod, K=1000, batch_size_K=100, noise=None):
        """multi_vaes.py:219-250: z ~ p(z)."""
        B = data[0].shape[0]
        L = self.params.latent_dim
        if noise is None:
            noise = draw_noise(self.params.dist, (B, K, L), data[0].device)
        loc, scale = self.pz_params[0].reshape(1, 1, L), self.pz_params[1].reshape(1, 1, L)
        if self.params.dist == "laplace":
            z = loc - scale * noise.sign() * torch.log1p(-noise.abs())
        else:
            z = loc + noise * scale
        z = z[:, :self._used(K, batch_size_K)].contiguous()
        ll = self._is_loglik(z, data, [mod], None, K_total=K, with_prior=False)
        return {f"uni_from_prior_{mod}": ll.mean()}

    @torch.no_grad()
    def compute_conditional_likelihood_bis(self, data, cond_mod, gen_mod, K=1000, batch_size_K=100, noise=None,
                                           noise_prior=None):
        """multi_vaes.py:253-268: ln p(x_gen|x_cond) ~ ln E_q(z|x_cond)[p(x,z)/q] - ln E_p(z)[p(x_cond|z)], each term by
        importance sampling.  The key carries the reference's trailing blank."""
        t1 = self.compute_joint_ll_from_uni(data, cond_mod, K, batch_size_K, noise=noise)[f"joint_ll_from_{cond_mod}"]
        t2 = self.compute_uni_ll_from_prior(data, cond_mod, K=K, batch_size_K=batch_size_K,
                                            noise=noise_prior)[f"uni_from_prior_{cond_mod}"]
        return {f"conditional_likelihood_bis_{cond_mod}_{gen_mod} ": t1 - t2}


class MMVAE(Multi_VAES):
    """mmvae.py:24-234."""

    def __init__(self, params, vaes):
        super().__init__(params, vaes)
        self.qz_x = dist_dict[params.dist]
        self.px_z_std = 1.0

    # ---- the fused core ----------------------------------------------------------------------------------
    def _encode_all(self, x):
        outs = self._run_branches([(lambda m=m, vae=vae: tuple(vae.encode_raw(x[m]))) for m, vae in enumerate(self.vaes)],
                                  shared=tuple(x))
        mus, raws = zip(*outs)
        return torch.stack(mus), torch.stack(raws)                                 # [M,B,L] each

    def _xform(self):
        return self.vaes[0].xform

    def _core(self, x, K, noise=None, cross=True, scaled=True):
        """-> mu [M,B,L], std [M,B,L], z [M,K,B,L], recon (list over decoders d of [E,K,B,*shape_d]),
        lpx [E(expert), M(decoder), K, B] (cross) or [M,K,B] (diagonal only)."""
        M = self.mod
        mu, raw = self._encode_all(x)
        _, B, L = mu.shape
        if noise is None:
            noise = draw_noise(self.params.dist, (M, K, B, L), mu.device)
        elif not torch.is_tensor(noise):
            noise = torch.stack(list(noise))
        z, std = ops.RSample.apply(mu, raw, noise.reshape(M, K, B, L), self.params.dist, self._xform(), False)
        def chain(d):
            zin = z.reshape(M * K * B, L) if cross else z[d].reshape(K * B, L)
            recon = self.vaes[d].decoder(zin)["reconstruction"]
            R = recon.shape[0]
            c = self.lik_scaling[d] if scaled else 1.0
            lp, lo = ops.LogLik.apply(recon.reshape(R, -1), self._flat_x(x[d]), 1, self.params.recon_losses[d], c, False)
            return recon, lp, lo

        recons, lps, los = [], [], []
        for recon, lp, lo in self._run_branches([(lambda d=d: chain(d)) for d in range(M)], shared=(z,) + tuple(x)):
            recons.append(recon.reshape(((M, K, B) if cross else (K, B)) + tuple(recon.shape[1:])))
            lps.append(lp.reshape((M, K, B) if cross else (K, B)))
            los.append(lo.reshape((M, K, B) if cross else (K, B)))
        lpx = torch.stack(lps, dim=1) if cross else torch.stack(lps, dim=0)
        self._lpx_lo = torch.stack(los, dim=1) if cross else torch.stack(los, dim=0)     # fp64 residuals of lpx
        return mu, std, z, recons, lpx

    def _likelihood(self, d, loc):
        if self.px_z[d] is dist.Bernoulli:
            return dist.Bernoulli(probs=loc, validate_args=False)
        return self.px_z[d](loc, torch.tensor(self.px_z_std, device=loc.device), validate_args=False)

    def forward(self, x, K=1, noise=None):
        """mmvae.py:31-80 -> (qz_xs, px_zs, zss, qz_x_params) with the reference shapes."""
        M = self.mod
        mu, std, z, recons, _ = self._core(x, K, noise)
        B, L = mu.shape[1:]
        qz_x_params = [(mu[m].unsqueeze(0).expand(K, B, L), std[m].unsqueeze(0).expand(K, B, L)) for m in range(M)]
        qz_xs = [self.qz_x(*p, validate_args=False) for p in qz_x_params]
        zss = [z[m] for m in range(M)]
        px_zs = [[self._likelihood(d, recons[d][e]) for d in range(M)] for e in range(M)]
        return qz_xs, px_zs, zss, qz_x_params

    def objective(self, x, K, mode, noise=None):
        """mode in iwae | iwae_looser | dreg | dreg_looser | elbo -> (scalar to MAXIMISE, aux dict)."""
        if mode == "elbo":
            mu, std, z, _, lpx = self._core(x, K, noise, cross=False)
            loss, terms, per = ops.Elbo.apply(z.detach(), mu, std, lpx, self.pz_params[0], self.pz_params[1],
                                              self.params.dist)
            return loss, {"terms": terms, "per_datapoint": per}
        mu, std, z, _, lpx = self._core(x, K, noise, cross=True)
        if mode in ("dreg", "dreg_looser"):                    # objectives.py:376,411: q-params detached in log q
            mu_q, std_q = mu.detach(), std.detach()
        else:
            mu_q, std_q = mu, std
        loss, lw, wt, per = ops.MoeLogW.apply(z, mu_q, std_q, lpx, self._lpx_lo, self.pz_params[0], self.pz_params[1],
                                              self.params.dist, mode)
        return loss, {"lw": lw, "grad_wt": wt, "per_datapoint": per}

    # ---- K-sample importance-sampled likelihoods ---------------------------------------------------------------
    def compute_qz_x_params(self, data):
        mu, raw = self._encode_all(data)
        M, B, L = mu.shape
        zero = torch.zeros((M, 1, B, L), device=mu.device)
        _, std = ops.RSample.apply(mu, raw, zero, "normal", self._xform(), False)
        return mu, std

    @torch.no_grad()
    def compute_joint_likelihood(self, data, K=1000, batch_size_K=100, mask=None, noise=None, correct_mixture=False):
        """mmvae.py:121-177.  mask (K,B) in {0,1} and noise [M] x (K,B,L) may be given for parity; the reference's
        `logsumexp(.)/2` mixture term (mmvae.py:166) is reproduced unless correct_mixture."""
        z, mu, std = self._joint_latents(data, K, batch_size_K, mask, noise)
        ll = self._is_loglik(z, data, range(self.mod), "mixture" if correct_mixture else "mixture_half", mu, std, K_total=K)
        self.last_per_datapoint = ll
        return {"likelihood": ll.mean()}

    def _joint_latents(self, data, K, batch_size_K, mask=None, noise=None):
        """mmvae.py:123-133: z_k from the 1/2-1/2 mixture of the two posteriors -> z [B,Kused,L], mu, std [M,B,L]."""
        assert self.mod == 2, "the reference's mixture sampling (mmvae.py:128-133) is written for two modalities"
        mu, raw = self._encode_all(data)
        M, B, L = mu.shape
        dev = mu.device
        if noise is None:
            noise = draw_noise(self.params.dist, (M, K, B, L), dev)
        elif not torch.is_tensor(noise):
            noise = torch.stack(list(noise))
        if mask is None:
            mask = torch.bernoulli(torch.full((K, B), 0.5, device=dev))
        zs, std = ops.RSample.apply(mu, raw, noise.reshape(M, K, B, L), self.params.dist, self._xform(), False)
        bern = mask.to(device=dev, dtype=torch.float32).reshape(K, B, 1)
        z = (bern * zs[0] + (1 - bern) * zs[1]).permute(1, 0, 2)[:, :self._used(K, batch_size_K)].contiguous()
        return z, mu, std

    @torch.no_grad()
    def compute_joint_ll_from_uni(self, data, cond_mod, K=1000, batch_size_K=100, noise=None):
        """mmvae.py:180-234: z ~ q(z|x_cond), weights p(x,y|z)p(z)/q(z|x_cond).  noise: (K,B,L)."""
        vae = self.vaes[cond_mod]
        mu, raw = vae.encode_raw(data[cond_mod])
        B, L = mu.shape
        if noise is None:
            noise = draw_noise(self.params.dist, (K, B, L), mu.device)
        z, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), noise.reshape(1, K, B, L), self.params.dist,
                                   vae.xform, False)
        z = z[0].permute(1, 0, 2)[:, :self._used(K, batch_size_K)].contiguous()
        ll = self._is_loglik(z, data, range(self.mod), "single", mu.unsqueeze(0), std, K_total=K)
        return {f"joint_ll_from_{cond_mod}": ll.mean()}


# ---------------------------------------------------------------------------------------- dataset subclasses
def _vae_class(params):
    return my_VAE if params.dist == "normal" else laplace_VAE


def _pixel_ratio_scaling(shapes, llik_scaling):
    """lik_scaling = (D_big / D_m) when params.llik_scaling == 0 (mmvae_mnist_svhn.py:53)."""
    sizes = [math.prod(s) for s in shapes]
    if llik_scaling == 0:
        big = max(sizes)
        return tuple(big / s for s in sizes)
    return tuple(float(llik_scaling) if s != max(sizes) else 1.0 for s in sizes)


class MNIST_SVHN(MMVAE):
    """mmvae_mnist_svhn.py:33-54: MNIST MLP enc/dec + SVHN conv enc/dec, lik_scaling = (3072/784, 1)."""
    shape_mods = [(1, 28, 28), (3, 32, 32)]

    def __init__(self, params):
        cfg1, cfg2 = VAEConfig((1, 28, 28), params.latent_dim), VAEConfig((3, 32, 32), params.latent_dim)
        vae = _vae_class(params)
        vaes = nn.ModuleList([vae(cfg1, Encoder_VAE_MLP(cfg1), Decoder_AE_MLP(cfg1)),
                              vae(cfg2, Encoder_VAE_SVHN(cfg2), Decoder_VAE_SVHN(cfg2))])
        super().__init__(params, vaes)
        self.modelName = "mmvae_mnist_svhn"
        self.vaes[0].modelName, self.vaes[1].modelName = "mnist", "svhn"
        s = getattr(params, "llik_scaling", 0.0)
        self.lik_scaling = ((3 * 32 * 32) / (1 * 28 * 28), 1.0) if s == 0 else (float(s), 1.0)


class CIRCLES_SQUARES(MMVAE):
    """mmvae_circles.py:33-59: default MLPs on (1,32,32), my_VAE/laplace_VAE, lik_scaling (1,1)."""
    shape_mods = [(1, 32, 32), (1, 32, 32)]

    def __init__(self, params):
        vae = _vae_class(params)
        cfgs = [VAEConfig((1, 32, 32), params.latent_dim) for _ in range(2)]
        vaes = nn.ModuleList([vae(c, Encoder_VAE_MLP(c), Decoder_AE_MLP(c)) for c in cfgs])
        super().__init__(params, vaes)
        self.modelName = "mmvae_circles_squares"
        self.vaes[0].modelName, self.vaes[1].modelName = "squares", "circles"
        self.lik_scaling = (1.0, 1.0)


class MNIST_SVHN_FASHION(MMVAE):
    """mmvae/mnist_svhn_fashion.py:30-53: trimodal, lik_scaling = (3072/784, 1, 3072/784)."""
    shape_mods = [(1, 28, 28), (3, 32, 32), (1, 28, 28)]

    def __init__(self, params):
        vae = _vae_class(params)
        c1, c2, c3 = (VAEConfig(s, params.latent_dim) for s in self.shape_mods)
        vaes = nn.ModuleList([vae(c1, Encoder_VAE_MLP(c1), Decoder_AE_MLP(c1)),
                              vae(c2, Encoder_VAE_SVHN(c2), Decoder_VAE_SVHN(c2)),
                              vae(c3, Encoder_VAE_MLP(c3), Decoder_AE_MLP(c3))])
        super().__init__(params, vaes)
        self.modelName = "mmvae_msf"
        for v, n in zip(self.vaes, ("mnist", "svhn", "fashion")):
            v.modelName = n
        r = (3 * 32 * 32) / (1 * 28 * 28)
        self.lik_scaling = (r, 1.0, r) if getattr(params, "llik_scaling", 0.0) == 0 else (1.0, 1.0, 1.0)


# registry names of models/__init__.py:2-33
VAE_mnist_svhn = MNIST_SVHN
VAE_circles_squares = CIRCLES_SQUARES
VAE_mmvae_msf = MNIST_SVHN_FASHION
MMVAE_CIRCLES = CIRCLES_SQUARES
