"""Encoder / decoder networks of the hot path, with the reference's class names, constructor arguments, output
containers and state_dict keys (SURVEY.md §8(b) "Checkpoint compatibility"), so reference checkpoints load.

torch.nn.Linear / Conv2d / ConvTranspose2d objects are used ONLY as parameter containers (same default init and
key names as the reference); their forward is never called — every contraction goes through libmmvae_b200
(ops.LinearAct / Conv2dAct / ConvT2dAct).

Reference: pythae 0.0.8 Encoder_VAE_MLP / Decoder_AE_MLP (restated; call sites mmvae_mnist_svhn.py:40-41),
models/nn/encoders.py:72-136 (SVHN conv stacks), :176-210 (TwoStepsEncoder), models/nn/joint_encoders.py:56-108.
"""
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from . import ops


class ModelOutput(OrderedDict):
    """pythae's ModelOutput: an OrderedDict with attribute access and integer indexing (joint_encoders.py:78 uses [0])."""

    def __getitem__(self, k):
        if isinstance(k, int):
            return list(self.values())[k]
        return super().__getitem__(k)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


class BaseEncoder(nn.Module):
    pass


class BaseDecoder(nn.Module):
    pass


def _linear(lin, x, act=None):
    return ops.LinearAct.apply(x, lin.weight, lin.bias, act)


def _flat_rows(x):
    return x.reshape(x.shape[0], -1)


class Encoder_VAE_MLP(BaseEncoder):
    """flatten -> Linear(D,512)+ReLU -> {embedding, log_var}: Linear(512,L)."""

    def __init__(self, args):
        super().__init__()
        self.input_dim = args.input_dim
        self.latent_dim = args.latent_dim
        self.layers = nn.ModuleList([nn.Sequential(nn.Linear(int(np.prod(args.input_dim)), 512), nn.ReLU())])
        self.depth = 1
        self.embedding = nn.Linear(512, self.latent_dim)
        self.log_var = nn.Linear(512, self.latent_dim)

    def forward(self, x, output_layer_levels=None):
        h = _linear(self.layers[0][0], _flat_rows(x), "relu")
        return ModelOutput(embedding=_linear(self.embedding, h), log_covariance=_linear(self.log_var, h))


class Decoder_AE_MLP(BaseDecoder):
    """Linear(L,512)+ReLU -> Linear(512,D)+Sigmoid -> reshape (.,C,H,W)."""

    def __init__(self, args):
        super().__init__()
        self.input_dim = args.input_dim
        self.latent_dim = args.latent_dim
        self.layers = nn.ModuleList([nn.Sequential(nn.Linear(args.latent_dim, 512), nn.ReLU()),
                                     nn.Sequential(nn.Linear(512, int(np.prod(args.input_dim))), nn.Sigmoid())])
        self.depth = 2

    def forward(self, z, output_layer_levels=None):
        lead = z.shape[:-1]
        h = _linear(self.layers[0][0], z.reshape(-1, z.shape[-1]), "relu")
        out = _linear(self.layers[1][0], h, "sigmoid")
        return ModelOutput(reconstruction=out.reshape(*lead, *self.input_dim))


class Encoder_VAE_SVHN(BaseEncoder):
    """models/nn/encoders.py:72-105."""

    def __init__(self, args):
        super().__init__()
        self.input_dim = args.input_dim
        self.latent_dim = args.latent_dim
        self.n_channels = args.input_dim[0]
        self.fBase = fB = 32
        self.enc = nn.Sequential(nn.Conv2d(self.n_channels, fB, 4, 2, 1, bias=True), nn.ReLU(True),
                                 nn.Conv2d(fB, fB * 2, 4, 2, 1, bias=True), nn.ReLU(True),
                                 nn.Conv2d(fB * 2, fB * 4, 4, 2, 1, bias=True), nn.ReLU(True))
        self.c1 = nn.Conv2d(fB * 4, self.latent_dim, 4, 2, 0)
        self.c2 = nn.Conv2d(fB * 4, self.latent_dim, 4, 2, 0)

    def forward(self, x):
        e = x
        for i in (0, 2, 4):
            e = ops.Conv2dAct.apply(e, self.enc[i].weight, self.enc[i].bias, 2, 1, "relu")
        mu = ops.Conv2dAct.apply(e, self.c1.weight, self.c1.bias, 2, 0, None).reshape(-1, self.latent_dim)
        lv = ops.Conv2dAct.apply(e, self.c2.weight, self.c2.bias, 2, 0, None).reshape(-1, self.latent_dim)
        return ModelOutput(embedding=mu, log_covariance=lv)


class Decoder_VAE_SVHN(BaseDecoder):
    """models/nn/encoders.py:108-136; accepts leading dims (:132-134)."""

    def __init__(self, args):
        super().__init__()
        self.latent_dim = args.latent_dim
        self.fBase = fB = 32
        self.nb_channels = args.input_dim[0]
        self.dec = nn.Sequential(nn.ConvTranspose2d(self.latent_dim, fB * 4, 4, 1, 0, bias=True), nn.ReLU(True),
                                 nn.ConvTranspose2d(fB * 4, fB * 2, 4, 2, 1, bias=True), nn.ReLU(True),
                                 nn.ConvTranspose2d(fB * 2, fB, 4, 2, 1, bias=True), nn.ReLU(True),
                                 nn.ConvTranspose2d(fB, self.nb_channels, 4, 2, 1, bias=True), nn.Sigmoid())

    def forward(self, z):
        lead = z.shape[:-1]
        h = z.reshape(-1, z.shape[-1], 1, 1)
        d = self.dec
        # every layer returns the gradient w.r.t. the PRE-activation of the layer below (its ReLU backward fused into the
        # dgrad), so layers 0, 2, 4 receive pre-activation gradients; h is never an autograd leaf here (z -> reshape).
        h = ops.ConvT2dAct.apply(h, d[0].weight, d[0].bias, 1, 0, "relu", False, True, False)
        h = ops.ConvT2dAct.apply(h, d[2].weight, d[2].bias, 2, 1, "relu", False, True, True)
        h = ops.ConvT2dAct.apply(h, d[4].weight, d[4].bias, 2, 1, "relu", False, True, True)
        out = ops.ConvT2dAct.apply(h, d[6].weight, d[6].bias, 2, 1, "sigmoid", True, False, True)  # NCHW at the boundary
        return ModelOutput(reconstruction=out.reshape(*lead, *out.shape[1:]))

    # Measured at C4 (500 datapoints, K = 1000): 7.84k datapoints/s fused against 8.05k unfused -- the per-image reduction
    # (a barrier among the four finalize warps, which also feed the MMA pipeline) costs more than the 12 KB store + re-read per
    # latent row it saves, so the fused kernel is opt-in.
    fuse_eval_loglik = False

    def loglik_rows(self, z, x_flat, x_div, lik, scale):
        """Evaluators (torch.no_grad): sum_pix log p(x | decoder(z_r)) for every latent row r WITHOUT materialising the
        reconstruction -- the last layer's kernel reduces the likelihood in its epilogue (mmvae.py:154-156 on top of
        encoders.py:125-128).  Returns None when the fused kernel does not apply; the caller then decodes and reduces."""
        d = self.dec
        if not self.fuse_eval_loglik or torch.is_grad_enabled() or not ops.convt_fwd_loglik_supported(d[6].weight, 16, 16, lik):
            return None
        h = z.reshape(-1, z.shape[-1], 1, 1)
        h = ops.ConvT2dAct.apply(h, d[0].weight, d[0].bias, 1, 0, "relu", False, True, False)
        h = ops.ConvT2dAct.apply(h, d[2].weight, d[2].bias, 2, 1, "relu", False, True, True)
        h = ops.ConvT2dAct.apply(h, d[4].weight, d[4].bias, 2, 1, "relu", False, True, True)
        return ops.convt_small_fwd_loglik(h, d[6].weight, d[6].bias, "sigmoid", x_flat, x_div, lik, scale)


class TwoStepsEncoder(BaseEncoder):
    """models/nn/encoders.py:176-210: frozen first stage under no_grad, then 3 x (Linear+ReLU) and two heads."""

    def __init__(self, first_encoder, args):
        super().__init__()
        self.input_dim = getattr(args, "input_dim", None)
        self.latent_dim = args.latent_dim
        self.first_encoder = first_encoder
        self.hidden_dim = 512
        self.num_hidden = 3
        mods = [nn.Linear(first_encoder.latent_dim, self.hidden_dim), nn.ReLU(True)]
        for _ in range(self.num_hidden - 1):
            mods += [nn.Linear(self.hidden_dim, self.hidden_dim), nn.ReLU(True)]
        self.sec_encoder = nn.Sequential(*mods)
        self.embedding = nn.Linear(self.hidden_dim, self.latent_dim)
        self.log_var = nn.Linear(self.hidden_dim, self.latent_dim)

    def forward(self, x):
        with torch.no_grad():
            h1 = self.first_encoder(x).embedding
        h = h1.detach()
        for i in range(0, 2 * self.num_hidden, 2):
            h = _linear(self.sec_encoder[i], h, "relu")
        return ModelOutput(embedding=_linear(self.embedding, h), log_covariance=_linear(self.log_var, h))


class LinearCCAEncoder(nn.Module):
    """Shape of the frozen DCCA first stage (dcca/models/mnist_svhn.py:50-104): encoder -> (h - m) @ w -> [:, :dim]."""

    def __init__(self, encoder, dcca_latent, dim):
        super().__init__()
        self.encoder = encoder
        self.latent_dim = dim
        self.register_buffer("m", torch.zeros(dcca_latent))
        self.register_buffer("w", torch.zeros(dcca_latent, dcca_latent))

    def forward(self, x):
        h = self.encoder(x).embedding
        # (h - m) @ w == h @ w - m @ w: one GEMM with bias = -(m @ w); w is stored [K, N] (b_kmajor = 0)
        R, Kd = h.shape
        bias = ops.gemm(self.m.reshape(1, Kd).contiguous(), self.w, None, 1, Kd, Kd, True, False).reshape(-1).neg()
        out = ops.LinearAct.apply(h, self.w.t().contiguous(), bias, None)       # differentiable w.r.t. h (DCCA-space
        return ModelOutput(embedding=out[:, :self.latent_dim].contiguous())    # reconstruction loss, jmvae_nf.py:147-156)


class DoubleHeadJoint(BaseEncoder):
    """models/nn/joint_encoders.py:56-82."""

    def __init__(self, hidden_dim, pretrained_config1, pretrained_config2, Encoder1, Encoder2, args):
        super().__init__()
        self.input1 = Encoder1(pretrained_config1)
        self.input2 = Encoder2(pretrained_config2)
        mods = [nn.Sequential(nn.Linear(pretrained_config1.latent_dim + pretrained_config2.latent_dim, hidden_dim),
                              nn.ReLU(True))]
        mods += [nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU(True))
                 for _ in range(args.num_hidden_layers - 1)]
        self.enc = nn.Sequential(*mods)
        self.fc21 = nn.Linear(hidden_dim, args.latent_dim)
        self.fc22 = nn.Linear(hidden_dim, args.latent_dim)

    def _heads(self, x):
        return [self.input1(x[0])[0], self.input2(x[1])[0]]

    def forward_raw(self, x):
        """-> (mu, raw) with std = exp(0.5 raw) + 1e-6 applied by the rsample kernel (xform 'exp_half_eps')."""
        e = torch.cat(self._heads(x), dim=1)
        for blk in self.enc:
            e = _linear(blk[0], e, "relu")
        return _linear(self.fc21, e), _linear(self.fc22, e)

    def forward(self, x):
        """joint_encoders.py:76-82: -> (mu, std) with std = exp(0.5*fc22(e)) + eta."""
        mu, raw = self.forward_raw(x)
        zero = torch.zeros((1, 1) + tuple(mu.shape), device=mu.device)
        _, std = ops.RSample.apply(mu.unsqueeze(0), raw.unsqueeze(0), zero, "normal", "exp_half_eps", False)
        return mu, std.reshape(mu.shape)


class MultipleHeadJoint(DoubleHeadJoint):
    """models/nn/joint_encoders.py:86-108."""

    def __init__(self, hidden_dim, pretrained_configs, encoders, args):
        BaseEncoder.__init__(self)
        self.inputs = nn.ModuleList([enc(cfg) for enc, cfg in zip(encoders, pretrained_configs)])
        mods = [nn.Sequential(nn.Linear(sum(c.latent_dim for c in pretrained_configs), hidden_dim), nn.ReLU(True))]
        mods += [nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU(True))
                 for _ in range(args.num_hidden_layers - 1)]
        self.enc = nn.Sequential(*mods)
        self.fc21 = nn.Linear(hidden_dim, args.latent_dim)
        self.fc22 = nn.Linear(hidden_dim, args.latent_dim)

    def _heads(self, x):
        return [enc(x[m])[0] for m, enc in enumerate(self.inputs)]
