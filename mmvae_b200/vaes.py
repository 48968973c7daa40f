"""Unimodal VAE wrappers with the reference's names and forward contract.

Reference: my_pythae/models/vae/vae_model_adapted.py:66-137 (my_VAE), laplace_vae/laplace_vae.py:51-91
(laplace_VAE), vae_iaf/vae_iaf_model_adapted.py:60-103, vae_maf/vae_maf_model_adapted.py:60-104.
`forward(x)` returns ModelOutput(recon_x, z0, mu, log_var, z, log_abs_det_jac, std).  The unused `neg_elbo` the
reference computes on every forward (vae_model_adapted.py:89) is not computed.
"""
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import ops
from .networks import Decoder_AE_MLP, Encoder_VAE_MLP, ModelOutput

F32_EPS = float(torch.finfo(torch.float32).eps)


@dataclass
class VAEConfig:
    input_dim: Optional[Tuple[int, ...]] = None
    latent_dim: int = 10
    uses_default_encoder: bool = True
    uses_default_decoder: bool = True
    reconstruction_loss: str = "mse"


@dataclass
class VAE_MAF_Config(VAEConfig):
    n_made_blocks: int = 2
    n_hidden_in_made: int = 3
    hidden_size: int = 128


@dataclass
class VAE_IAF_Config(VAE_MAF_Config):
    pass


def draw_noise(dist, shape, device, generator=None):
    """The noise torch.distributions would draw: Laplace.rsample uses u~U(eps-1, 1); Normal uses N(0,1)."""
    if dist == "laplace":
        return torch.empty(shape, device=device, dtype=torch.float32).uniform_(F32_EPS - 1.0, 1.0, generator=generator)
    return torch.randn(shape, device=device, dtype=torch.float32, generator=generator)


class _BaseVAE(nn.Module):
    dist = "normal"          # family of q(z0|x)
    xform = "exp_half"       # how the second encoder head becomes std

    def __init__(self, model_config, encoder=None, decoder=None):
        super().__init__()
        self.model_config = model_config
        self.input_dim = model_config.input_dim
        self.latent_dim = model_config.latent_dim
        # registration order decoder -> encoder, as pythae's BaseAE does
        self.decoder = decoder if decoder is not None else Decoder_AE_MLP(model_config)
        self.encoder = encoder if encoder is not None else Encoder_VAE_MLP(model_config)

    def set_encoder(self, encoder):
        self.encoder = encoder

    def set_decoder(self, decoder):
        self.decoder = decoder

    # -- pieces used by the fused multimodal paths
    def encode_raw(self, x):
        o = self.encoder(x)
        return o.embedding, o.log_covariance

    def flow_inverse(self, z0):
        return z0, None

    def flow_forward(self, z):
        return z, None

    def sample(self, x, noise=None, decode=True):
        mu, log_var = self.encode_raw(x)
        B, L = mu.shape
        if noise is None:
            noise = draw_noise(self.dist, (B, L), mu.device)
        z0, std = ops.RSample.apply(mu.unsqueeze(0), log_var.unsqueeze(0), noise.reshape(1, 1, B, L), self.dist,
                                    self.xform, False)
        z0, std = z0.reshape(B, L), std.reshape(B, L)
        z, ladj = self.flow_inverse(z0)
        if ladj is None:
            ladj = torch.zeros(1, device=mu.device)
        recon = self.decoder(z)["reconstruction"] if decode else None
        return ModelOutput(recon_x=recon, z0=z0, mu=mu, log_var=log_var, z=z, log_abs_det_jac=ladj, std=std)

    def forward(self, x, noise=None, **kwargs):
        return self.sample(x, noise=noise, decode=True)

    def inverse_flow(self, z):
        """vae_model_adapted.py:126-131: identity flow (JMVAE `no_nf`)."""
        out, ladj = self.flow_forward(z)
        if ladj is None:
            ladj = torch.zeros(z.shape[0], device=z.device)
        return ModelOutput(out=out, log_abs_det_jac=ladj)


class my_VAE(_BaseVAE):
    """Normal posterior, std = exp(0.5 log_var), z = mu + eps*std (vae_model_adapted.py:84-86,133-137)."""
    dist, xform = "normal", "exp_half"


class _FlowVAE(my_VAE):
    """Gaussian q(z0|x) pushed through a flow: z = flow.inverse(z0) (sampling direction), z0 = flow(z)."""
    flow_cls = None

    def __init__(self, model_config, encoder=None, decoder=None):
        super().__init__(model_config, encoder, decoder)
        from . import flows
        cls = getattr(flows, self.flow_cls)
        self.flow = cls(model_config.latent_dim, getattr(model_config, "n_made_blocks", 2),
                        getattr(model_config, "n_hidden_in_made", 3), getattr(model_config, "hidden_size", 128))

    def flow_inverse(self, z0):
        o = self.flow.inverse(z0)
        return o.out, o.log_abs_det_jac

    def flow_forward(self, z):
        o = self.flow(z)
        return o.out, o.log_abs_det_jac


class my_VAE_MAF(_FlowVAE):
    """vae_maf/vae_maf_model_adapted.py:39-104 (the unused neg_elbo of :90 is not computed — SURVEY A5)."""
    flow_cls = "MAF"


class my_VAE_IAF(_FlowVAE):
    """vae_iaf/vae_iaf_model_adapted.py:39-103."""
    flow_cls = "IAF"


class laplace_VAE(_BaseVAE):
    """Laplace posterior, std = softmax(log_var)*L + 1e-6 (laplace_vae.py:69-71)."""
    dist, xform = "laplace", "softmax"
