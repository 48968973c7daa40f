"""TEST INFRASTRUCTURE — CPU restatement ("port") of the multimodal-ELBO hot path of clementchadebec/mmvae
(`bivae`).  Plain torch-CPU tensor arithmetic, functional over a reference-keyed `state_dict`, with every
random draw taken as an explicit noise argument so the same z reaches the oracle, the reference and the
CUDA path.  Works in fp32 or fp64 (convert the state dict and inputs).

ONLY `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py` may
import this module; `mmvae_b200/` never does (the product path has no CPU fallback).

PARITY PIN: `tests/golden/*.pt` were produced by running the UNMODIFIED reference through
`oracle/ref_shim.py` in the build container (`tests/golden/make_golden.py`); `tests/test_oracle_golden.py`
checks this file against them (and, where `/root/reference` exists, against the live reference).
Caveat: `Encoder_VAE_MLP` / `Decoder_AE_MLP` come from pythae==0.0.8 (requirements.txt:11), which is not under
/root/reference; they are restated from upstream (layout confirmed by /root/reference/model2.pt) — that one
definition is "parity unpinned" by any reference-side test.

Citations are to /root/reference/src/bivae/… unless noted.
"""
import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

ETA = 1e-6                       # utils.py:17 Constants.eta
LOG2PI = math.log(2 * math.pi)


# ----------------------------------------------------------------------------------------------------------
# model description
# ----------------------------------------------------------------------------------------------------------
@dataclass
class Spec:
    """Architecture description of one reference model class (mmvae_mnist_svhn.py:33-54 etc.)."""
    shapes: List[Tuple[int, int, int]]
    enc: List[str]                      # 'mlp' | 'svhn' | 'twostep'
    dec: List[str]                      # 'mlp' | 'svhn'
    latent_dim: int
    dist: str = "laplace"               # posterior/prior family: 'laplace' | 'normal'
    recon_losses: Sequence[str] = ("normal", "normal")
    lik_scaling: Sequence[float] = (1.0, 1.0)
    vae: str = "laplace_VAE"            # 'laplace_VAE' | 'my_VAE' | 'my_VAE_MAF' | 'my_VAE_IAF'
    n_made_blocks: int = 2
    n_hidden_in_made: int = 3
    hidden_size: int = 128
    joint: Optional[str] = None         # None | 'double_head' | 'multi_head'
    joint_heads: List[str] = field(default_factory=list)   # per-modality head encoder type
    joint_head_latent: int = 20
    num_hidden_layers: int = 1
    first_enc: List[str] = field(default_factory=list)     # twostep: type of the frozen first encoder
    dcca_latent: int = 16
    dim_dcca: int = 9

    @property
    def M(self):
        return len(self.shapes)


def spec_mnist_svhn(latent_dim=20, dist="laplace", recon_losses=("normal", "normal"), llik_scaling=0.0):
    """mmvae_mnist_svhn.py:33-54."""
    s = (3 * 32 * 32) / (1 * 28 * 28) if llik_scaling == 0 else llik_scaling
    return Spec(shapes=[(1, 28, 28), (3, 32, 32)], enc=["mlp", "svhn"], dec=["mlp", "svhn"],
                latent_dim=latent_dim, dist=dist, recon_losses=tuple(recon_losses), lik_scaling=(s, 1.0),
                vae="my_VAE" if dist == "normal" else "laplace_VAE")


def spec_circles(latent_dim=2, dist="normal", recon_losses=("normal", "normal")):
    """mmvae_circles.py:33-59: default pythae MLPs on (1,32,32), my_VAE, lik_scaling (1,1)."""
    return Spec(shapes=[(1, 32, 32), (1, 32, 32)], enc=["mlp", "mlp"], dec=["mlp", "mlp"],
                latent_dim=latent_dim, dist=dist, recon_losses=tuple(recon_losses), lik_scaling=(1.0, 1.0),
                vae="my_VAE")


def spec_msf(latent_dim=30, dist="normal", recon_losses=("normal",) * 3, llik_scaling=0.0):
    """mmvae/mnist_svhn_fashion.py:30-53."""
    s = (3 * 32 * 32) / (1 * 28 * 28)
    ls = (s, 1.0, s) if llik_scaling == 0 else (1.0, 1.0, 1.0)
    return Spec(shapes=[(1, 28, 28), (3, 32, 32), (1, 28, 28)], enc=["mlp", "svhn", "mlp"],
                dec=["mlp", "svhn", "mlp"], latent_dim=latent_dim, dist=dist,
                recon_losses=tuple(recon_losses), lik_scaling=ls,
                vae="my_VAE" if dist == "normal" else "laplace_VAE")


def spec_jnf_mnist_svhn(latent_dim=20, flow="maf", no_nf=False, dcca=True, dim_dcca=9, n_made_blocks=2,
                        num_hidden_layers=1, llik_scaling=0.0, recon_losses=("normal", "normal")):
    """jmvae_nf_mnist_svhn_dcca.py:37-94."""
    s = (3 * 32 * 32) / (1 * 28 * 28) if llik_scaling == 0 else llik_scaling
    vae = "my_VAE" if no_nf else ("my_VAE_IAF" if flow == "iaf" else "my_VAE_MAF")
    return Spec(shapes=[(1, 28, 28), (3, 32, 32)], enc=["twostep", "twostep"] if dcca else ["mlp", "svhn"],
                dec=["mlp", "svhn"], latent_dim=latent_dim, dist="normal", recon_losses=tuple(recon_losses),
                lik_scaling=(s, 1.0), vae=vae, n_made_blocks=n_made_blocks, joint="double_head",
                joint_heads=["mlp", "svhn"], joint_head_latent=20, num_hidden_layers=num_hidden_layers,
                first_enc=["mlp", "svhn"], dcca_latent=16, dim_dcca=dim_dcca)


# ----------------------------------------------------------------------------------------------------------
# networks (L0)
# ----------------------------------------------------------------------------------------------------------
# Test hook for branch-consistent comparisons.  The gradient of a ReLU network is ill-conditioned at the kinks: a unit whose
# pre-activation is within the forward error of zero lands on either side, and each such flip moves the gradient by an O(1)
# term (at C2 size ~1e-6 of 10^8 units flip between ANY two fp32 evaluations, ~1e-3 of a gradient's norm).  With RELU_MASKS =
# {weight key: bool tensor [rows, ...]} the oracle takes the sign pattern of the layer fed by that weight from the
# implementation under test (rows consumed in call order), so what is compared is the arithmetic, not the coin flips.
RELU_MASKS = None
_relu_cursor = {}


def relu_masks(masks):
    """Install (or clear with None) the sign patterns; resets the per-key row cursors."""
    global RELU_MASKS
    RELU_MASKS = masks
    _relu_cursor.clear()


def _relu(x, key):
    if RELU_MASKS is None or key not in RELU_MASKS:
        return F.relu(x)
    m = RELU_MASKS[key]
    c = _relu_cursor.get(key, 0)
    n = x.shape[0]
    assert c + n <= m.shape[0] and tuple(m.shape[1:]) == tuple(x.shape[1:]), (key, c, n, tuple(m.shape), tuple(x.shape))
    _relu_cursor[key] = c + n
    return x * m[c:c + n].to(x.dtype)


def enc_mlp(sd, p, x):
    """pythae 0.0.8 Encoder_VAE_MLP: flatten → Linear(D,512)+ReLU → {embedding, log_var}: Linear(512,L)."""
    h = x.reshape(x.shape[0], -1)
    h = _relu(F.linear(h, sd[p + "layers.0.0.weight"], sd[p + "layers.0.0.bias"]), p + "layers.0.0.weight")
    return (F.linear(h, sd[p + "embedding.weight"], sd[p + "embedding.bias"]),
            F.linear(h, sd[p + "log_var.weight"], sd[p + "log_var.bias"]))


def dec_mlp(sd, p, z, shape):
    """pythae 0.0.8 Decoder_AE_MLP: Linear(L,512)+ReLU → Linear(512,D)+Sigmoid → reshape (·,C,H,W)."""
    h = _relu(F.linear(z, sd[p + "layers.0.0.weight"], sd[p + "layers.0.0.bias"]), p + "layers.0.0.weight")
    h = torch.sigmoid(F.linear(h, sd[p + "layers.1.0.weight"], sd[p + "layers.1.0.bias"]))
    return h.reshape((z.shape[0],) + tuple(shape))


def enc_svhn(sd, p, x, L):
    """models/nn/encoders.py:72-105."""
    e = x
    for i in (0, 2, 4):
        e = _relu(F.conv2d(e, sd[p + f"enc.{i}.weight"], sd[p + f"enc.{i}.bias"], stride=2, padding=1), p + f"enc.{i}.weight")
    mu = F.conv2d(e, sd[p + "c1.weight"], sd[p + "c1.bias"], stride=2, padding=0).reshape(-1, L)
    lv = F.conv2d(e, sd[p + "c2.weight"], sd[p + "c2.bias"], stride=2, padding=0).reshape(-1, L)
    return mu, lv


def dec_svhn(sd, p, z):
    """models/nn/encoders.py:108-136 (leading dims flattened by the caller)."""
    h = z.reshape(-1, z.shape[-1], 1, 1)
    h = _relu(F.conv_transpose2d(h, sd[p + "dec.0.weight"], sd[p + "dec.0.bias"], stride=1, padding=0), p + "dec.0.weight")
    h = _relu(F.conv_transpose2d(h, sd[p + "dec.2.weight"], sd[p + "dec.2.bias"], stride=2, padding=1), p + "dec.2.weight")
    h = _relu(F.conv_transpose2d(h, sd[p + "dec.4.weight"], sd[p + "dec.4.bias"], stride=2, padding=1), p + "dec.4.weight")
    h = torch.sigmoid(F.conv_transpose2d(h, sd[p + "dec.6.weight"], sd[p + "dec.6.bias"], stride=2, padding=1))
    return h


def enc_twostep(sd, p, x, first_kind, dcca_latent, dim_dcca):
    """models/nn/encoders.py:176-210 around a DCCA first stage of the shape of dcca/models/mnist_svhn.py:50-104:
    frozen encoder → (h − m) @ w → [:, :dim] (no grad), then 3×(Linear+ReLU) and two heads."""
    with torch.no_grad():
        fp = p + "first_encoder."
        if first_kind == "mlp":
            h, _ = enc_mlp(sd, fp + "encoder.", x)
        else:
            h, _ = enc_svhn(sd, fp + "encoder.", x, dcca_latent)
        h = torch.matmul(h - sd[fp + "m"], sd[fp + "w"])[:, :dim_dcca]
    for i in (0, 2, 4):
        h = _relu(F.linear(h, sd[p + f"sec_encoder.{i}.weight"], sd[p + f"sec_encoder.{i}.bias"]), p + f"sec_encoder.{i}.weight")
    return (F.linear(h, sd[p + "embedding.weight"], sd[p + "embedding.bias"]),
            F.linear(h, sd[p + "log_var.weight"], sd[p + "log_var.bias"]))


def encode(spec: Spec, sd, m, x):
    p = f"vaes.{m}.encoder."
    k = spec.enc[m]
    if k == "mlp":
        return enc_mlp(sd, p, x)
    if k == "svhn":
        return enc_svhn(sd, p, x, spec.latent_dim)
    if k == "twostep":
        return enc_twostep(sd, p, x, spec.first_enc[m], spec.dcca_latent, spec.dim_dcca)
    raise ValueError(k)


def decode(spec: Spec, sd, d, z):
    """z: (R, L) → (R, C, H, W)."""
    p = f"vaes.{d}.decoder."
    if spec.dec[d] == "mlp":
        return dec_mlp(sd, p, z, spec.shapes[d])
    return dec_svhn(sd, p, z)


def joint_encode(spec: Spec, sd, x):
    """joint_encoders.py:56-82 (DoubleHeadJoint) / :86-108 (MultipleHeadJoint) → (mu, std)."""
    hs = []
    for m in range(spec.M):
        p = (f"joint_encoder.input{m + 1}." if spec.joint == "double_head" else f"joint_encoder.inputs.{m}.")
        if spec.joint_heads[m] == "mlp":
            hs.append(enc_mlp(sd, p, x[m])[0])
        else:
            hs.append(enc_svhn(sd, p, x[m], spec.joint_head_latent)[0])
    e = torch.cat(hs, dim=1)
    for j in range(spec.num_hidden_layers):
        e = _relu(F.linear(e, sd[f"joint_encoder.enc.{j}.0.weight"], sd[f"joint_encoder.enc.{j}.0.bias"]), f"joint_encoder.enc.{j}.0.weight")
    mu = F.linear(e, sd["joint_encoder.fc21.weight"], sd["joint_encoder.fc21.bias"])
    std = torch.exp(0.5 * F.linear(e, sd["joint_encoder.fc22.weight"], sd["joint_encoder.fc22.bias"])) + ETA
    return mu, std


# ----------------------------------------------------------------------------------------------------------
# MADE / MAF / IAF  (my_pythae/models/normalizing_flows/…)
# ----------------------------------------------------------------------------------------------------------
def made_masks(L, hidden, n_hidden):
    """made/made_model.py:68-95, sequential ordering."""
    m = {-1: torch.arange(L)}
    for i in range(n_hidden):
        m[i] = torch.arange(hidden) % (L - 1)
    masks = [(m[i].unsqueeze(-1) >= m[i - 1].unsqueeze(0)).float() for i in range(n_hidden)]
    masks.append((m[n_hidden - 1].unsqueeze(0) < m[-1].unsqueeze(-1)).float().repeat(2, 1))
    return masks


def made(sd, p, x, n_hidden):
    """made_model.py:97-111 + layers.py:24-25: F.linear(x, mask*W, b), ReLU between, → (mu, log_var)."""
    h = x
    for i in range(n_hidden + 1):
        w = sd[p + f"net.{2 * i}.mask"].to(h.dtype) * sd[p + f"net.{2 * i}.weight"]
        h = F.linear(h, w, sd[p + f"net.{2 * i}.bias"])
        if i < n_hidden:
            h = _relu(h, p + f"net.{2 * i}.bias")
    L = x.shape[1]
    return h[:, :L], h[:, L:]


def maf_forward(spec: Spec, sd, p, x):
    """maf/maf_model.py:49-76 (data → prior): one MADE pass per block."""
    ladj = x.new_zeros(x.shape[0])
    for b in range(spec.n_made_blocks):
        mu, lv = made(sd, p + f"net.{b}.", x, spec.n_hidden_in_made)
        x = (x - mu) * (-lv).exp()
        ladj = ladj - lv.sum(-1)
        x = x.flip(dims=(1,))
    return x, ladj


def maf_inverse(spec: Spec, sd, p, y):
    """maf/maf_model.py:78-110 (prior → data): L sequential MADE passes per block, blocks reversed."""
    ladj = y.new_zeros(y.shape[0])
    L = y.shape[1]
    for b in reversed(range(spec.n_made_blocks)):
        y = y.flip(dims=(1,))
        cols = []
        x = torch.zeros_like(y)
        for i in range(L):
            mu, lv = made(sd, p + f"net.{b}.", x, spec.n_hidden_in_made)
            xi = y[:, i] * lv[:, i].exp() + mu[:, i]
            cols.append(xi)
            x = torch.cat([torch.stack(cols, 1), torch.zeros_like(y[:, i + 1:])], 1)
            ladj = ladj + lv[:, i]
        y = x
    return y, ladj


def iaf_forward(spec: Spec, sd, p, x):
    """iaf/iaf_model.py:48-80: the sequential direction."""
    ladj = x.new_zeros(x.shape[0])
    L = x.shape[1]
    for b in range(spec.n_made_blocks):
        cols = []
        y = torch.zeros_like(x)
        for i in range(L):
            m, s = made(sd, p + f"net.{b}.", y, spec.n_hidden_in_made)
            cols.append((x[:, i] - m[:, i]) * (-s[:, i]).exp())
            y = torch.cat([torch.stack(cols, 1), torch.zeros_like(x[:, i + 1:])], 1)
            ladj = ladj - s[:, i]
        x = y.flip(dims=(1,))
    return x, ladj


def iaf_inverse(spec: Spec, sd, p, y):
    """iaf/iaf_model.py:82-107: one pass per block, blocks reversed."""
    ladj = y.new_zeros(y.shape[0])
    for b in reversed(range(spec.n_made_blocks)):
        y = y.flip(dims=(1,))
        m, s = made(sd, p + f"net.{b}.", y, spec.n_hidden_in_made)
        y = y * s.exp() + m
        ladj = ladj + s.sum(-1)
    return y, ladj


def flow_forward(spec, sd, m, z):
    """vae.flow(z) if the VAE has a flow, else inverse_flow identity (jmvae_nf.py:64; vae_model_adapted.py:126-131)."""
    p = f"vaes.{m}.flow."
    if spec.vae == "my_VAE_MAF":
        return maf_forward(spec, sd, p, z)
    if spec.vae == "my_VAE_IAF":
        return iaf_forward(spec, sd, p, z)
    return z, z.new_zeros(z.shape[0])


def flow_inverse(spec, sd, m, z0):
    p = f"vaes.{m}.flow."
    if spec.vae == "my_VAE_MAF":
        return maf_inverse(spec, sd, p, z0)
    if spec.vae == "my_VAE_IAF":
        return iaf_inverse(spec, sd, p, z0)
    return z0, z0.new_zeros(z0.shape[0])


# ----------------------------------------------------------------------------------------------------------
# distributions as the reference uses them (torch.distributions)
# ----------------------------------------------------------------------------------------------------------
def lat_log_prob(dist, z, loc, scale):
    """Laplace/Normal(loc, scale).log_prob(z), elementwise."""
    if dist == "laplace":
        return -torch.log(2 * scale) - torch.abs(z - loc) / scale
    return -((z - loc) ** 2) / (2 * scale ** 2) - torch.log(scale) - math.log(math.sqrt(2 * math.pi))


def lat_rsample(dist, loc, scale, noise):
    """torch Laplace.rsample: loc − scale·sign(u)·log1p(−|u|), u~U(eps−1,1); Normal: loc + eps·scale."""
    if dist == "laplace":
        return loc - scale * noise.sign() * torch.log1p(-noise.abs())
    return loc + noise * scale


def lat_kl(dist, loc, scale, ploc, pscale):
    """torch.distributions.kl: _kl_laplace_laplace / _kl_normal_normal."""
    if dist == "laplace":
        r = scale / pscale
        d = (loc - ploc).abs()
        return -r.log() + d / pscale + r * torch.exp(-d / scale) - 1
    vr = (scale / pscale).pow(2)
    t1 = ((loc - ploc) / pscale).pow(2)
    return 0.5 * (vr + t1 - 1 - vr.log())


def px_log_prob(kind, recon, x):
    """px_z(recon, scale=1).log_prob(x) elementwise (mmvae.py:55-61; Bernoulli(probs=recon))."""
    if kind == "normal":
        return -((x - recon) ** 2) / 2 - math.log(math.sqrt(2 * math.pi))
    if kind == "laplace":
        return -math.log(2.0) - torch.abs(x - recon)
    if kind == "bernoulli":
        eps = torch.finfo(recon.dtype).eps
        p = recon.clamp(min=eps, max=1 - eps)
        logits = torch.log(p) - torch.log1p(-p)
        return -F.binary_cross_entropy_with_logits(logits, x.expand_as(logits), reduction="none")
    raise ValueError(kind)


def posterior_std(spec: Spec, log_var):
    """laplace_vae.py:69  /  vae_model_adapted.py:84."""
    if spec.vae == "laplace_VAE":
        return F.softmax(log_var, dim=-1) * log_var.size(-1) + 1e-6
    return torch.exp(0.5 * log_var)


def pz_params(spec: Spec, sd, like):
    if "_pz_params.0" in sd:
        return sd["_pz_params.0"].to(like.dtype), sd["_pz_params.1"].to(like.dtype)
    return like.new_zeros(1, spec.latent_dim), like.new_ones(1, spec.latent_dim)


# ----------------------------------------------------------------------------------------------------------
# MMVAE.forward (mmvae.py:31-80) — K-tiling removed (results-neutral)
# ----------------------------------------------------------------------------------------------------------
def mmvae_forward(spec: Spec, sd, x, K, noise, cross=True):
    """noise[m]: (K,B,L) — u~U(eps−1,1) for 'laplace', eps~N(0,1) for 'normal' (row k·B+b of the reference's
    tiled batch).  Returns mus[m] (B,L), stds[m] (B,L), zss[m] (K,B,L), recon[e][d] (K,B,C,H,W)."""
    M = spec.M
    mus, stds, zss = [], [], []
    for m in range(M):
        mu, lv = encode(spec, sd, m, x[m])
        std = posterior_std(spec, lv)
        z = lat_rsample(spec.dist, mu.unsqueeze(0), std.unsqueeze(0), noise[m])
        mus.append(mu), stds.append(std), zss.append(z)
    B = x[0].shape[0]
    L = spec.latent_dim
    recon = [[None] * M for _ in range(M)]
    for e in range(M):
        for d in range(M):
            if e == d or cross:
                r = decode(spec, sd, d, zss[e].reshape(K * B, L))
                recon[e][d] = r.reshape((K, B) + tuple(spec.shapes[d]))
    return mus, stds, zss, recon


def _lpx(spec, recon_ed, x_d, d, scaled=True):
    """Σ_pix log p(x_d|z)·lik_scaling[d]  → (K,B)   (objectives.py:381-383)."""
    lp = px_log_prob(spec.recon_losses[d], recon_ed, x_d)
    lp = lp.reshape(lp.shape[0], lp.shape[1], -1)
    if scaled:
        lp = lp * spec.lik_scaling[d]
    return lp.sum(-1)


def _log_mean_exp(v, dim=0):
    return torch.logsumexp(v, dim) - math.log(v.size(dim))          # utils.py:143-144


def mmvae_lws(spec: Spec, sd, x, K, noise, detach_q):
    """lw[r] (K,B) for each expert r  (objectives.py:117-131 / 372-389 / 406-422)."""
    mus, stds, zss, recon = mmvae_forward(spec, sd, x, K, noise)
    ploc, pscale = pz_params(spec, sd, mus[0])
    qm = [(m.detach(), s.detach()) if detach_q else (m, s) for m, s in zip(mus, stds)]
    lws = []
    for r in range(spec.M):
        lpz = lat_log_prob(spec.dist, zss[r], ploc, pscale).sum(-1)
        lq = torch.stack([lat_log_prob(spec.dist, zss[r], m.unsqueeze(0), s.unsqueeze(0)).sum(-1)
                          for (m, s) in qm])
        lqz_x = _log_mean_exp(lq)
        lpx_z = torch.stack([_lpx(spec, recon[r][d], x[d], d) for d in range(spec.M)]).sum(0)
        lws.append(lpz + lpx_z - lqz_x)
    return lws, (mus, stds, zss, recon)


def m_iwae(spec, sd, x, K, noise):
    """objectives.py:333-340 with the 4-tuple unpack widened."""
    lws, _ = mmvae_lws(spec, sd, x, K, noise, detach_q=False)
    return _log_mean_exp(torch.cat(lws)).sum()


def m_iwae_looser(spec, sd, x, K, noise):
    """objectives.py:361-369."""
    lws, _ = mmvae_lws(spec, sd, x, K, noise, detach_q=False)
    return _log_mean_exp(torch.stack(lws), dim=1).mean(0).sum()


def m_dreg(spec, sd, x, K, noise):
    """objectives.py:391-403; the zss hook never fires → weights are constants."""
    lws, _ = mmvae_lws(spec, sd, x, K, noise, detach_q=True)
    lw = torch.cat(lws)
    with torch.no_grad():
        grad_wt = (lw - torch.logsumexp(lw, 0, keepdim=True)).exp()
    return (grad_wt * lw).sum()


def m_dreg_looser(spec, sd, x, K, noise):
    """objectives.py:425-438."""
    lws, _ = mmvae_lws(spec, sd, x, K, noise, detach_q=True)
    lw = torch.stack(lws)
    with torch.no_grad():
        grad_wt = (lw - torch.logsumexp(lw, 1, keepdim=True)).exp()
    return (grad_wt * lw).mean(0).sum()


def m_elbo(spec, sd, x, K, noise):
    """objectives.py:87-111: closed-form KL + importance ratio on the DIAGONAL decoders only."""
    mus, stds, zss, recon = mmvae_forward(spec, sd, x, K, noise, cross=False)
    ploc, pscale = pz_params(spec, sd, mus[0])
    M = spec.M
    lpx_zs, klds = [], []
    for r in range(M):
        kld = lat_kl(spec.dist, mus[r], stds[r], ploc, pscale).sum(-1)          # (B,) — K identical copies
        klds.append(kld.unsqueeze(0).expand(K, -1))
        for d in range(M):
            lpx_z = _lpx(spec, recon[d][d], x[d], d)
            if d == r:
                lwt = lpx_z.new_zeros(())
            else:
                zs = zss[d].detach()
                lwt = (lat_log_prob(spec.dist, zs, mus[r].unsqueeze(0), stds[r].unsqueeze(0))
                       - lat_log_prob(spec.dist, zs, mus[d].unsqueeze(0), stds[d].unsqueeze(0)).detach()).sum(-1)
            lpx_zs.append(lwt.exp() * lpx_z)
    obj = (1 / M) * (torch.stack(lpx_zs).sum(0) - torch.stack(klds).sum(0))
    details = {f"lpx_zs{i // 2}{i % 2}": lpx_zs[i].sum() for i in range(4)}
    return obj.mean(0).sum(), details


OBJECTIVES = {"elbo": lambda *a: m_elbo(*a)[0], "iwae": m_iwae, "iwae_looser": m_iwae_looser,
              "dreg": m_dreg, "dreg_looser": m_dreg_looser}


# ----------------------------------------------------------------------------------------------------------
# unimodal VAE forward (L1) — what `vaes[m](x)` returns that the evaluators use
# ----------------------------------------------------------------------------------------------------------
def vae_sample(spec: Spec, sd, m, x, noise):
    """(mu, log_var, std, z0, z, ladj) of `vaes[m](x)`; decoder not run.
    laplace_vae.py:65-71; vae_model_adapted.py:80-86; vae_maf_model_adapted.py:77-87; vae_iaf_…:77-87."""
    mu, lv = encode(spec, sd, m, x)
    std = posterior_std(spec, lv)
    z0 = lat_rsample("laplace" if spec.vae == "laplace_VAE" else "normal", mu, std, noise)
    z, ladj = flow_inverse(spec, sd, m, z0)
    return mu, lv, std, z0, z, ladj


def _lpx_unscaled(spec, d, recon, x_d):
    """Σ_{c,h,w} px_z(recon, scale=1).log_prob(x_d)  (mmvae.py:154-156) — no lik_scaling in evaluators (A7)."""
    lp = px_log_prob(spec.recon_losses[d], recon, x_d)
    return lp.reshape(lp.shape[0], lp.shape[1], -1).sum(-1)


def _decode_all(spec, sd, z_bkl, data):
    """Σ_m Σ_pix log p(x_m[b] | dec_m(z[b,k]))  for z (B,Kc,L) → (B,Kc)."""
    B, Kc, L = z_bkl.shape
    tot = 0
    for m in range(spec.M):
        r = decode(spec, sd, m, z_bkl.reshape(B * Kc, L)).reshape((B, Kc) + tuple(spec.shapes[m]))
        tot = tot + _lpx_unscaled(spec, m, r, data[m].unsqueeze(1))
    return tot


def _used(K, batch_size_K):
    """`while stop_index <= K` drops the tail chunk but ln K is still subtracted."""
    return (K // batch_size_K) * batch_size_K


def mmvae_joint_likelihood(spec, sd, data, K, batch_size_K, mask, noise, correct_mixture=False):
    """MMVAE.compute_joint_likelihood, mmvae.py:121-177.  mask: (K,B) in {0,1} (np.random.binomial, :128-130);
    noise[m]: (K,B,L) for q_m.rsample([K]) (:133).  Returns (mean, per-datapoint (B,))."""
    assert spec.M == 2
    qp = []
    for m in range(2):
        mu, lv = encode(spec, sd, m, data[m])
        qp.append((mu, posterior_std(spec, lv)))
    bern = mask.to(qp[0][0].dtype).unsqueeze(-1)
    zs = [lat_rsample(spec.dist, qp[m][0].unsqueeze(0), qp[m][1].unsqueeze(0), noise[m]) for m in range(2)]
    z = (bern * zs[0] + (1 - bern) * zs[1]).permute(1, 0, 2)               # (B,K,L)
    z = z[:, :_used(K, batch_size_K)]
    ploc, pscale = pz_params(spec, sd, z)
    lpx = _decode_all(spec, sd, z, data)
    lpz = lat_log_prob(spec.dist, z, ploc, pscale).sum(-1)
    lq = torch.stack([lat_log_prob(spec.dist, z, mu.unsqueeze(1), std.unsqueeze(1)).sum(-1) for mu, std in qp])
    lqz = torch.logsumexp(lq, 0) - math.log(2) if correct_mixture else torch.logsumexp(lq, 0) / 2   # :166 (A8)
    ll = torch.logsumexp(lpx + lpz - lqz, dim=1) - math.log(K)
    return ll.mean(), ll


def mmvae_joint_ll_from_uni(spec, sd, data, cond_mod, K, batch_size_K, noise):
    """MMVAE.compute_joint_ll_from_uni, mmvae.py:180-234.  noise: (K,B,L)."""
    mu, lv = encode(spec, sd, cond_mod, data[cond_mod])
    std = posterior_std(spec, lv)
    z = lat_rsample(spec.dist, mu.unsqueeze(0), std.unsqueeze(0), noise).permute(1, 0, 2)
    z = z[:, :_used(K, batch_size_K)]
    ploc, pscale = pz_params(spec, sd, z)
    a = (_decode_all(spec, sd, z, data) + lat_log_prob(spec.dist, z, ploc, pscale).sum(-1)
         - lat_log_prob(spec.dist, z, mu.unsqueeze(1), std.unsqueeze(1)).sum(-1))
    ll = torch.logsumexp(a, dim=1) - math.log(K)
    return ll.mean(), ll


def conditional_likelihood(spec, sd, data, cond_mod, gen_mod, K, batch_size_K, noise):
    """Multi_VAES.compute_conditional_likelihood, multi_vaes.py:271-318.  noise: (B,Kused,L) — the draw for
    datapoint i, chunk c, row j sits at [i, c·batch_size_K + j].  Returns (mean, per-datapoint (B,))."""
    B = data[0].shape[0]
    Ku = _used(K, batch_size_K)
    L = spec.latent_dim
    xr = data[cond_mod].unsqueeze(1).expand((B, Ku) + tuple(data[cond_mod].shape[1:]))
    xr = xr.reshape((B * Ku,) + tuple(data[cond_mod].shape[1:]))
    _, _, _, _, z, _ = vae_sample(spec, sd, cond_mod, xr, noise.reshape(B * Ku, L))
    r = decode(spec, sd, gen_mod, z).reshape((B, Ku) + tuple(spec.shapes[gen_mod]))
    lp = _lpx_unscaled(spec, gen_mod, r, data[gen_mod].unsqueeze(1))
    ll = torch.logsumexp(lp, dim=1) - math.log(K)
    return ll.sum() / B, ll


def conditional_likelihoods(spec, sd, data, K, batch_size_K, noises):
    """Multi_VAES.compute_conditional_likelihoods, multi_vaes.py:324-355.  noises[(cond,gen)]."""
    M = spec.M
    metrics, ll = {}, [[None] * M for _ in range(M)]
    for i in range(M):
        for j in range(M):
            if i != j:
                mean, per = conditional_likelihood(spec, sd, data, j, i, K, batch_size_K, noises[(j, i)])
                metrics[f"cond_likelihood_{j}_{i}"] = mean
                ll[i][j] = per
    if M == 3:
        for i in range(3):
            moe = torch.logsumexp(torch.stack([ll[i][j] for j in range(M) if i != j]), dim=0) - math.log(2)
            metrics[f"cond_lw_subset_{i}"] = moe.mean()
    return metrics, ll


def uni_ll_from_prior(spec, sd, data, mod, K, batch_size_K, noise):
    """Multi_VAES.compute_uni_ll_from_prior, multi_vaes.py:219-250.  noise: (B,K,L) prior rsample noise."""
    ploc, pscale = pz_params(spec, sd, noise)
    z = lat_rsample(spec.dist, ploc, pscale, noise)[:, :_used(K, batch_size_K)]
    B, Kc, L = z.shape
    r = decode(spec, sd, mod, z.reshape(B * Kc, L)).reshape((B, Kc) + tuple(spec.shapes[mod]))
    ll = torch.logsumexp(_lpx_unscaled(spec, mod, r, data[mod].unsqueeze(1)), dim=1) - math.log(K)
    return ll.mean(), ll


# ----------------------------------------------------------------------------------------------------------
# JMVAE-NF (jmvae_nf.py, objectives.py:177-220)
# ----------------------------------------------------------------------------------------------------------
def _recon_loss(kind, recon, x):
    """recon_loss_dict, objectives.py:177 — 'sum' of the elementwise loss (MSE has no ½)."""
    r, t = recon.reshape(x.shape[0], -1), x.reshape(x.shape[0], -1)
    if kind == "normal":
        return F.mse_loss(r, t, reduction="none").sum()
    if kind == "laplace":
        return F.l1_loss(r, t, reduction="none").sum()
    return F.binary_cross_entropy(r, t, reduction="none").sum()


def jmvae_nf_compute_kld(spec, sd, x, eps2, eps_uni=None, beta_kl=1.0, no_recon=True):
    """JMVAE_NF.compute_kld, jmvae_nf.py:56-85 (second stochastic forward, its decodes discarded — A14)."""
    mu, std = joint_encode(spec, sd, x)
    z = mu + eps2 * std
    lq_xy = lat_log_prob("normal", z, mu, std).sum()
    reg, det = 0, {}
    for m in range(spec.M):
        z0, ladj = flow_forward(spec, sd, m, z)
        mu_m, lv_m = encode(spec, sd, m, x[m])
        log_q_z0 = (-0.5 * (lv_m + LOG2PI + torch.pow(z0 - mu_m, 2) / torch.exp(lv_m))).sum(dim=1)
        det[f"kld_{m}"] = lq_xy - (log_q_z0 + ladj).sum()
        if no_recon:
            reg = reg + beta_kl * det[f"kld_{m}"]
        else:
            _, _, _, _, zz, _ = vae_sample(spec, sd, m, x[m], eps_uni[m])
            rec = decode(spec, sd, m, zz)
            det[f"recon_loss_{m}"] = F.mse_loss(x[m].reshape(x[m].shape[0], -1),
                                                rec.reshape(x[m].shape[0], -1), reduction="sum")
            reg = reg + beta_kl * det[f"kld_{m}"] + det[f"recon_loss_{m}"]
    return reg, det


def m_jmvae_nf(spec, sd, x, eps1, eps2, eps_uni=None, epoch=1, warmup=0, beta_prior=1.0, beta_kl=1.0,
               no_recon=True, linear_warmup=False):
    """objectives.py:179-220.  Freezing of joint encoder/decoders (:180-183) is the caller's business:
    pass state-dict tensors with the matching requires_grad."""
    beta_reg = ((epoch - 1) / warmup if epoch - 1 <= warmup else 1) if linear_warmup else 1
    mu, std = joint_encode(spec, sd, x)
    z = mu + eps1 * std
    loss, details = 0, {}
    for m in range(spec.M):
        rec = decode(spec, sd, m, z)
        details[f"loss_{m}"] = _recon_loss(spec.recon_losses[m], rec, x[m]) * spec.lik_scaling[m]
        loss = loss - details[f"loss_{m}"]
    details["loss"] = loss
    lv = 2 * torch.log(std)
    details["kld_prior"] = -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp(), dim=-1).sum()
    if epoch >= warmup or linear_warmup:
        reg, det = jmvae_nf_compute_kld(spec, sd, x, eps2, eps_uni, beta_kl, no_recon)
        details["reg"] = reg
        details.update(det)
    else:
        details["reg"] = 0
    return loss - beta_reg * (beta_prior * details["kld_prior"] + details["reg"]), details


def jmvae_nf_joint_likelihood(spec, sd, data, K, batch_size_K, noise):
    """JMVAE_NF.compute_joint_likelihood, jmvae_nf.py:209-270.  noise: (K,B,L)."""
    mu, std = joint_encode(spec, sd, data)
    z = (mu.unsqueeze(0) + noise * std.unsqueeze(0)).permute(1, 0, 2)[:, :_used(K, batch_size_K)]
    ploc, pscale = pz_params(spec, sd, z)
    a = (_decode_all(spec, sd, z, data) + lat_log_prob("normal", z, ploc, pscale).sum(-1)
         - lat_log_prob("normal", z, mu.unsqueeze(1), std.unsqueeze(1)).sum(-1))
    ll = torch.logsumexp(a, dim=1) - math.log(K)
    return ll.mean(), ll


def jmvae_nf_joint_ll_from_uni(spec, sd, data, cond_mod, K, batch_size_K, noise):
    """JMVAE_NF.compute_joint_ll_from_uni, jmvae_nf.py:87-143.  noise: (B,Kused,L)."""
    B = data[0].shape[0]
    Ku = _used(K, batch_size_K)
    L = spec.latent_dim
    xc = data[cond_mod]
    xr = xc.unsqueeze(1).expand((B, Ku) + tuple(xc.shape[1:])).reshape((B * Ku,) + tuple(xc.shape[1:]))
    mu, lv, _, z0, z, ladj = vae_sample(spec, sd, cond_mod, xr, noise.reshape(B * Ku, L))
    log_q_z0 = (-0.5 * (lv + LOG2PI + torch.pow(z0 - mu, 2) / torch.exp(lv))).sum(dim=1)
    lqz_x = (log_q_z0 - ladj).reshape(B, Ku)
    zb = z.reshape(B, Ku, L)
    ploc, pscale = pz_params(spec, sd, zb)
    a = _decode_all(spec, sd, zb, data) + lat_log_prob("normal", zb, ploc, pscale).sum(-1) - lqz_x
    ll = torch.logsumexp(a, dim=1) - math.log(K)
    return ll.mean(), ll


# ----------------------------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------------------------
def compute_microbatch_split(x, K):
    """objectives.py:10-17."""
    B = x[0].size(0)
    S = sum(1.0 / (K * math.prod(_x.size()[1:])) for _x in x)
    S = int(1e8 * S)
    assert S > 0, "Cannot fit individual data in memory, consider smaller K"
    return min(B, S)


def logsumexp_combine(ms, ss):
    """Stable merge of per-shard (max, Σexp) pairs: the reference's own two-level form (mmvae.py:168-175)."""
    m = torch.stack(ms).max(0).values
    s = sum(s_i * torch.exp(m_i - m) for m_i, s_i in zip(ms, ss))
    return m + torch.log(s)


# ----------------------------------------------------------------------------------------------------------
# MVAE (product of experts with the prior) — models/mvae/mvae.py; objective `m_self_built` (objectives.py:481-483)
# ----------------------------------------------------------------------------------------------------------
def mvae_poe(mus, log_vars):
    """MVAE.poe, mvae.py:27-45: product of the Gaussian experts and the N(0,1) prior -> (joint_mu, joint_std)."""
    mus = list(mus) + [torch.zeros_like(mus[0])]
    log_vars = list(log_vars) + [torch.zeros_like(log_vars[0])]
    lnT = torch.stack([-l for l in log_vars])
    lnV = -torch.logsumexp(lnT, dim=0)
    joint_mu = (torch.exp(lnT) * torch.stack(mus)).sum(dim=0) * torch.exp(lnV)
    return joint_mu, torch.exp(0.5 * lnV)


def mvae_forward(spec: Spec, sd, x, noise):
    """MVAE.forward, mvae.py:74-143 without subsampling.  noise: M+1 tensors (B,L) in the order the reference draws them
    (one per unimodal PoE, then the joint).  -> dict(elbo, z_joint, joint_mu, joint_std)."""
    ploc, pscale = None, None
    elbo = 0
    mus, lvs = [], []
    for m in range(spec.M):
        mu_u, lv_u = encode(spec, sd, m, x[m])
        mus.append(mu_u)
        lvs.append(lv_u)
        mu, std = mvae_poe([mu_u], [lv_u])
        if ploc is None:
            ploc, pscale = pz_params(spec, sd, mu)
        z = mu + noise[m] * std
        recon = decode(spec, sd, m, z)
        elbo = elbo - 0.5 * torch.sum((x[m] - recon) ** 2) * spec.lik_scaling[m] - lat_kl("normal", mu, std, ploc, pscale).sum()
    joint_mu, joint_std = mvae_poe(mus, lvs)
    z_joint = joint_mu + noise[spec.M] * joint_std
    for m in range(spec.M):
        recon = decode(spec, sd, m, z_joint)
        elbo = elbo - 0.5 * torch.sum((x[m] - recon) ** 2) * spec.lik_scaling[m]
    elbo = elbo - lat_kl("normal", joint_mu, joint_std, ploc, pscale).sum()
    return {"elbo": elbo, "z_joint": z_joint, "joint_mu": joint_mu, "joint_std": joint_std}


def mvae_joint_likelihood(spec, sd, data, K, batch_size_K, noise):
    """MVAE.compute_joint_likelihood, mvae.py:219-266: z ~ N(joint_mu, joint_std), K // batch_size_K chunks of
    batch_size_K draws per datapoint, ln K subtracted.  noise: (B, Kused, L).  -> (mean, per-datapoint (B,))."""
    enc = [encode(spec, sd, m, data[m]) for m in range(spec.M)]
    mu, std = mvae_poe([e[0] for e in enc], [e[1] for e in enc])
    z = mu.unsqueeze(1) + noise * std.unsqueeze(1)                              # (B,Ku,L)
    ploc, pscale = pz_params(spec, sd, z)
    a = (_decode_all(spec, sd, z, data) + lat_log_prob("normal", z, ploc, pscale).sum(-1)
         - lat_log_prob("normal", z, mu.unsqueeze(1), std.unsqueeze(1)).sum(-1))
    ll = torch.logsumexp(a, dim=1) - math.log(K)
    return ll.mean(), ll


def mvae_joint_ll_from_uni(spec, sd, data, cond_mod, K, batch_size_K, noise):
    """MVAE.compute_joint_ll_from_uni, mvae.py:160-217: z ~ q(z|x_cond) = the ENCODER's Gaussian (not the PoE with the
    prior — the reference's own warning at :172-173).  noise: (K,B,L) (`rsample([K])`)."""
    mu, lv = encode(spec, sd, cond_mod, data[cond_mod])
    std = torch.exp(0.5 * lv)
    z = (mu.unsqueeze(0) + noise * std.unsqueeze(0)).permute(1, 0, 2)[:, :_used(K, batch_size_K)]
    ploc, pscale = pz_params(spec, sd, z)
    a = (_decode_all(spec, sd, z, data) + lat_log_prob("normal", z, ploc, pscale).sum(-1)
         - lat_log_prob("normal", z, mu.unsqueeze(1), std.unsqueeze(1)).sum(-1))
    ll = torch.logsumexp(a, dim=1) - math.log(K)
    return ll.mean(), ll


# ----------------------------------------------------------------------------------------------------------
# MoPoE (mixture of products of experts) — models/moepoe/moepoe.py
# ----------------------------------------------------------------------------------------------------------
def mopoe_components(mus, log_vars):
    """The mixture components of MOEPOE.forward (moepoe.py:93-100): the M unimodal posteriors followed by the PoE of
    every subset of >= 2 modalities in itertools.combinations order; only the FULL subset includes the N(0,1) prior
    (compute_poe_for_all_subsets, :44-70).  -> (list of mu, list of log-variance)."""
    from itertools import combinations
    M = len(mus)
    out_mu, out_lv = list(mus), list(log_vars)
    for k in range(2, M + 1):
        for tup in combinations(range(M), k):
            ms = [mus[i] for i in tup]
            ls = [log_vars[i] for i in tup]
            if k == M:
                ms, ls = ms + [torch.zeros_like(mus[0])], ls + [torch.zeros_like(log_vars[0])]
            lnT = torch.stack([-l for l in ls])
            lnV = -torch.logsumexp(lnT, dim=0)
            out_mu.append((torch.exp(lnT) * torch.stack(ms)).sum(dim=0) * torch.exp(lnV))
            out_lv.append(lnV)
    return out_mu, out_lv


def mopoe_select(mus, log_vars):
    """mixture_component_selection, moepoe.py:20-40: datapoints [k*(B//C), (k+1)*(B//C)) take component k, the last
    component takes the remainder."""
    C, B = len(mus), mus[0].shape[0]
    step = B // C
    bounds = [(k * step, B if k == C - 1 else (k + 1) * step) for k in range(C)]
    return (torch.cat([mus[k][a:b] for k, (a, b) in enumerate(bounds)]),
            torch.cat([log_vars[k][a:b] for k, (a, b) in enumerate(bounds)]))


def mopoe_forward(spec: Spec, sd, x, noise, beta_kl=1.0):
    """MOEPOE.forward, moepoe.py:83-135.  noise: (B,L).  -> dict(elbo, z_joint, mus, log_vars)."""
    enc = [encode(spec, sd, m, x[m]) for m in range(spec.M)]
    mus, lvs = mopoe_components([e[0] for e in enc], [e[1] for e in enc])
    mu_r, lv_r = mopoe_select(mus, lvs)
    z = mu_r + noise * torch.exp(0.5 * lv_r)
    ploc, pscale = pz_params(spec, sd, z)
    elbo = 0
    for m in range(spec.M):
        recon = decode(spec, sd, m, z)
        elbo = elbo + px_log_prob(spec.recon_losses[m], recon, x[m]).sum() * spec.lik_scaling[m]
    for mu, lv in zip(mus, lvs):
        elbo = elbo - lat_kl("normal", mu, torch.exp(0.5 * lv), ploc, pscale).sum() * beta_kl / len(mus)
    return {"elbo": elbo, "z_joint": z, "mus": torch.stack(mus), "log_vars": torch.stack(lvs)}


def mopoe_joint_ll_from_uni(spec, sd, data, cond_mod, K, batch_size_K, noise):
    """MOEPOE.compute_joint_ll_from_uni, moepoe.py:160-214: as the MVAE one but ln K is NOT subtracted (:212)."""
    mean, ll = mvae_joint_ll_from_uni(spec, sd, data, cond_mod, K, batch_size_K, noise)
    return mean + math.log(K), ll + math.log(K)


# ----------------------------------------------------------------------------------------------------------
# MMVAE-NF (mixture of flow-transformed unimodal posteriors) — models/mmvae_nf/mmvae_nf.py; objective `m_elbo_nf`
# ----------------------------------------------------------------------------------------------------------
def spec_mmvae_nf_mnist_svhn(latent_dim=20, n_made_blocks=2, llik_scaling=0.0):
    """mmvae_nf/mnist_svhn.py:41-68: my_VAE_IAF unimodal VAEs (pythae MLP / SVHN conv nets), Gaussian prior."""
    s = (3 * 32 * 32) / (1 * 28 * 28) if llik_scaling == 0 else llik_scaling
    return Spec(shapes=[(1, 28, 28), (3, 32, 32)], enc=["mlp", "svhn"], dec=["mlp", "svhn"], latent_dim=latent_dim,
                dist="normal", recon_losses=("normal", "normal"), lik_scaling=(s, 1.0), vae="my_VAE_IAF",
                n_made_blocks=n_made_blocks)


def mmvae_nf_forward(spec: Spec, sd, x, noise):
    """MMVAE_NF.forward, mmvae_nf.py:30-62 -> (ln_qz_xs [M x (B,)], zs [M x (B,L)], recons[e][d]).
    ln q(z_m|x_m) = sum_L -1/2 (lv + (z0-mu)^2 / e^lv) - log|det J| (no 2 pi term, :47-50).  noise: M tensors (B,L)."""
    ln_q, zs = [], []
    for m in range(spec.M):
        mu, lv, _, z0, z, ladj = vae_sample(spec, sd, m, x[m], noise[m])
        ln_q.append((-0.5 * (lv + torch.pow(z0 - mu, 2) / torch.exp(lv))).sum(dim=1) - ladj)
        zs.append(z)
    recons = [[decode(spec, sd, d, zs[e]) for d in range(spec.M)] for e in range(spec.M)]
    return ln_q, zs, recons


def m_elbo_nf(spec: Spec, sd, x, noise):
    """objectives.py:463-479.  `row[e].sum()` (:471) indexes the (B,) tensor ln q(z_e|x_e) with the MODALITY index e, so
    only datapoint e's log-density enters the KL term — the reference's arithmetic, reproduced as is."""
    ln_q, zs, recons = mmvae_nf_forward(spec, sd, x, noise)
    elbo = 0
    for e in range(spec.M):
        kld = ln_q[e][e].sum() - (-0.5 * torch.sum(zs[e] ** 2))
        elbo = elbo - kld / spec.M
        for d in range(spec.M):
            elbo = elbo + (-0.5 * torch.sum((recons[e][d] - x[d]) ** 2)) / spec.M * spec.lik_scaling[d]
    return elbo


# ----------------------------------------------------------------------------------------------------------
# product-of-experts SUBSET posteriors: density, Hamiltonian sampler, importance-sampled conditional likelihood
# (jmvae_nf.py:272-437; mvae.py:268-327; modalities/trimodal.py:158-230)
# ----------------------------------------------------------------------------------------------------------
def jmvae_nf_poe_posterior(spec: Spec, sd, subset, z, data, divide_prior=True):
    """JMVAE_NF.compute_poe_posterior value, jmvae_nf.py:294-329: sum_{m in subset} [log N(flow_m(z); mu_m, e^{lv_m}) +
    log|det J_m|] - log p(z) (when divide_prior) -> (n,)."""
    lnq = 0
    if divide_prior:
        lnq = lnq + (0.5 * (torch.pow(z, 2) + LOG2PI)).sum(dim=1)
    for m in subset:
        z0, ladj = flow_forward(spec, sd, m, z)
        mu, lv = encode(spec, sd, m, data[m])
        lnq = lnq + (-0.5 * (lv + LOG2PI + torch.pow(z0 - mu, 2) / torch.exp(lv))).sum(dim=1) + ladj
    return lnq


def jmvae_nf_poe_posterior_grad(spec: Spec, sd, subset, z, data, divide_prior=True):
    """(value, d value.sum() / dz) as jmvae_nf.py:307,323-327."""
    zz = z.clone().detach().requires_grad_(True)
    with torch.enable_grad():
        v = jmvae_nf_poe_posterior(spec, sd, subset, zz, data, divide_prior)
        g = torch.autograd.grad(v.sum(), zz)[0]
    return v.detach(), g


def jmvae_nf_hmc_poe_subset(spec: Spec, sd, subset, data, K, z_init, gammas, accepts, n_lf=10, eps_lf=0.01,
                            divide_prior=True):
    """JMVAE_NF.sample_from_poe_subset, jmvae_nf.py:332-437.  data is repeated K times (row k*n+b); z_init (K*n,L) is
    the mixture-of-experts start (sample_from_moe_subset), gammas [steps,(K*n,L)] the momenta, accepts [steps,(K*n,)]
    the uniforms of the accept test.  -> (z [K,n,L], acceptance counts (K*n,))."""
    n = len(data[0])
    data = [torch.cat([d] * K) for d in data]
    z0 = z_init
    z = z0
    acc_nbr = torch.zeros(len(z))
    for gamma, u in zip(gammas, accepts):
        rho = gamma
        lnq, g = jmvae_nf_poe_posterior_grad(spec, sd, subset, z, data, divide_prior)
        H0 = -lnq + 0.5 * torch.norm(rho, dim=1) ** 2
        for _ in range(n_lf):
            rho_ = rho - (eps_lf / 2) * (-g)
            z = z + eps_lf * rho_
            lnq, g = jmvae_nf_poe_posterior_grad(spec, sd, subset, z, data, divide_prior)
            rho = rho_ - (eps_lf / 2) * (-g)
        H = -lnq + 0.5 * torch.norm(rho, dim=1) ** 2
        alpha = torch.exp(H0 - H)
        moves = (u < alpha).to(z.dtype).reshape(-1, 1)
        acc_nbr = acc_nbr + moves[:, 0]
        z = z * moves + (1 - moves) * z0
        z0 = z
    return z.reshape(K, n, -1), acc_nbr


def mvae_poe_posterior(spec: Spec, sd, subset, z, data):
    """MVAE.compute_poe_posterior, mvae.py:304-327.  `poe` returns (mu, STD) and the density below reads the second value
    as a log-variance — the reference's own arithmetic, reproduced as is."""
    lq = 0
    for m in subset:
        mu_u, lv_u = encode(spec, sd, m, data[m])
        t0, t1 = mvae_poe([mu_u], [lv_u])
        lq = lq + (-0.5 * (t1 + LOG2PI + torch.pow(z - t0, 2) / torch.exp(t1))).sum(dim=1)
    return lq - (-0.5 * (LOG2PI + torch.pow(z, 2))).sum(dim=1)


def cond_ll_from_poe_subset(kind, spec: Spec, sd, data, subset, gen_mod, K, batch_size_K, i_s, noise):
    """compute_cond_ll_from_poe_subset, modalities/trimodal.py:158-216.  kind: 'mvae' | 'jmvae_nf'.  i_s (B,) in {0,1}: the
    member of the subset used as the importance distribution for each datapoint (np.random.randint, :184); noise
    (B, nb*batch_size_K, L).  As in the reference, `lnpxs` is re-created inside the chunk loop (:188), so only the LAST
    chunk enters the numerator while every chunk enters the normaliser.  -> (mean, per-datapoint (B,))."""
    nb = K // batch_size_K
    out = []
    for i in range(len(data[0])):
        a, o = subset[int(i_s[i])], subset[1 - int(i_s[i])]
        norm, last = [], None
        for n in range(nb):
            eps = noise[i, n * batch_size_K:(n + 1) * batch_size_K]
            rep = [torch.stack(batch_size_K * [data[m][i]]) for m in range(spec.M)]
            if kind == "mvae":
                mu_u, lv_u = encode(spec, sd, a, rep[a])
                mu, std = mvae_poe([mu_u], [lv_u])
                z = mu + eps * std
                lnq = mvae_poe_posterior(spec, sd, [o], z, rep)
            else:
                z = vae_sample(spec, sd, a, rep[a], eps)[4]
                lnq = jmvae_nf_poe_posterior(spec, sd, [o], z, rep, True)
            recon = decode(spec, sd, gen_mod, z)
            lpx = px_log_prob(spec.recon_losses[gen_mod], recon, data[gen_mod][i]).sum(dim=(1, 2, 3)) + lnq
            norm.append(torch.logsumexp(lnq, dim=0))
            last = torch.logsumexp(lpx, dim=0)
        out.append(last - torch.logsumexp(torch.stack(norm), dim=0))
    per = torch.stack(out)
    return per.sum() / len(per), per
