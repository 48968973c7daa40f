"""TEST INFRASTRUCTURE — deterministic synthetic weights / inputs / noise, built on numpy's frozen legacy
`RandomState` stream so fixtures can store a seed instead of megabytes of weights and still be regenerated
bit-identically on any box.  Shapes follow the reference's loaders: MNIST (B,1,28,28),
SVHN (B,3,32,32), circles/squares (B,1,32,32), fp32 in [0,1].
"""
import math

import numpy as np
import torch

from . import bivae_oracle as O

F32_EPS = float(torch.finfo(torch.float32).eps)


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a.astype(np.float32)))


def _fan_in(shape, transposed=False):
    if len(shape) == 2:
        return shape[1]
    if len(shape) == 4:                       # Conv (Cout,Cin,kh,kw) fan_in=Cin·kh·kw; ConvT (Cin,Cout,kh,kw): torch uses size(1)·kh·kw
        return shape[1] * shape[2] * shape[3]
    return shape[0]


def _linear(rs, sd, p, out_f, in_f):
    b = 1.0 / math.sqrt(in_f)
    sd[p + "weight"] = _t(rs.uniform(-b, b, (out_f, in_f)))
    sd[p + "bias"] = _t(rs.uniform(-b, b, (out_f,)))


def _conv(rs, sd, p, shape):
    b = 1.0 / math.sqrt(_fan_in(shape))
    sd[p + "weight"] = _t(rs.uniform(-b, b, shape))
    nb = shape[0]
    sd[p + "bias"] = _t(rs.uniform(-b, b, (nb,)))


def _convT(rs, sd, p, shape):
    b = 1.0 / math.sqrt(_fan_in(shape))
    sd[p + "weight"] = _t(rs.uniform(-b, b, shape))
    sd[p + "bias"] = _t(rs.uniform(-b, b, (shape[1],)))


def _enc_mlp(rs, sd, p, D, L):
    _linear(rs, sd, p + "layers.0.0.", 512, D)
    _linear(rs, sd, p + "embedding.", L, 512)
    _linear(rs, sd, p + "log_var.", L, 512)


def _dec_mlp(rs, sd, p, D, L):
    _linear(rs, sd, p + "layers.0.0.", 512, L)
    _linear(rs, sd, p + "layers.1.0.", D, 512)


def _enc_svhn(rs, sd, p, C, L):
    _conv(rs, sd, p + "enc.0.", (32, C, 4, 4))
    _conv(rs, sd, p + "enc.2.", (64, 32, 4, 4))
    _conv(rs, sd, p + "enc.4.", (128, 64, 4, 4))
    _conv(rs, sd, p + "c1.", (L, 128, 4, 4))
    _conv(rs, sd, p + "c2.", (L, 128, 4, 4))


def _dec_svhn(rs, sd, p, C, L):
    _convT(rs, sd, p + "dec.0.", (L, 128, 4, 4))
    _convT(rs, sd, p + "dec.2.", (128, 64, 4, 4))
    _convT(rs, sd, p + "dec.4.", (64, 32, 4, 4))
    _convT(rs, sd, p + "dec.6.", (32, C, 4, 4))


def _flow(rs, sd, p, spec):
    L, H, nh = spec.latent_dim, spec.hidden_size, spec.n_hidden_in_made
    masks = O.made_masks(L, H, nh)
    sizes = [L] + [H] * nh + [2 * L]
    for b in range(spec.n_made_blocks):
        for i in range(nh + 1):
            q = p + f"net.{b}.net.{2 * i}."
            _linear(rs, sd, q, sizes[i + 1], sizes[i])
            sd[q + "mask"] = masks[i].clone()


def make_state_dict(spec: O.Spec, seed: int, weight_gain: float = 1.0):
    """Reference-keyed state dict (SURVEY §8(b) 'Checkpoint compatibility') with torch-default-like scales.
    Insertion order mirrors the reference modules' registration order (decoder before encoder: pythae BaseAE)."""
    rs = np.random.RandomState(seed)
    sd = {}
    L = spec.latent_dim
    if spec.joint is not None:
        for m in range(spec.M):
            p = f"joint_encoder.input{m + 1}." if spec.joint == "double_head" else f"joint_encoder.inputs.{m}."
            C, H, W = spec.shapes[m]
            if spec.joint_heads[m] == "mlp":
                _enc_mlp(rs, sd, p, C * H * W, spec.joint_head_latent)
            else:
                _enc_svhn(rs, sd, p, C, spec.joint_head_latent)
        _linear(rs, sd, "joint_encoder.enc.0.0.", 512, spec.joint_head_latent * spec.M)
        for j in range(1, spec.num_hidden_layers):
            _linear(rs, sd, f"joint_encoder.enc.{j}.0.", 512, 512)
        _linear(rs, sd, "joint_encoder.fc21.", L, 512)
        _linear(rs, sd, "joint_encoder.fc22.", L, 512)
    for m in range(spec.M):
        C, H, W = spec.shapes[m]
        D = C * H * W
        p = f"vaes.{m}."
        if spec.dec[m] == "mlp":
            _dec_mlp(rs, sd, p + "decoder.", D, L)
        else:
            _dec_svhn(rs, sd, p + "decoder.", C, L)
        if spec.enc[m] == "mlp":
            _enc_mlp(rs, sd, p + "encoder.", D, L)
        elif spec.enc[m] == "svhn":
            _enc_svhn(rs, sd, p + "encoder.", C, L)
        else:
            fp = p + "encoder.first_encoder."
            if spec.first_enc[m] == "mlp":
                _enc_mlp(rs, sd, fp + "encoder.", D, spec.dcca_latent)
            else:
                _enc_svhn(rs, sd, fp + "encoder.", C, spec.dcca_latent)
            sd[fp + "m"] = _t(rs.uniform(-0.1, 0.1, (spec.dcca_latent,)))
            sd[fp + "w"] = _t(rs.uniform(-0.5, 0.5, (spec.dcca_latent, spec.dcca_latent)))
            _linear(rs, sd, p + "encoder.sec_encoder.0.", 512, spec.dim_dcca)
            _linear(rs, sd, p + "encoder.sec_encoder.2.", 512, 512)
            _linear(rs, sd, p + "encoder.sec_encoder.4.", 512, 512)
            _linear(rs, sd, p + "encoder.embedding.", L, 512)
            _linear(rs, sd, p + "encoder.log_var.", L, 512)
        if spec.vae in ("my_VAE_MAF", "my_VAE_IAF"):
            _flow(rs, sd, p + "flow.", spec)
    sd["_pz_params.0"] = torch.zeros(1, L)
    sd["_pz_params.1"] = torch.ones(1, L)
    if weight_gain != 1.0:
        for k in sd:
            if k.endswith("weight"):
                sd[k] = sd[k] * weight_gain
    return sd


def make_inputs(spec: O.Spec, B: int, seed: int, binary=()):
    """U(0,1) images of the reference loader shapes; modalities listed in `binary` are {0,1}
    (Bernoulli.log_prob raises on non-binary targets)."""
    rs = np.random.RandomState(seed)
    xs = []
    for m, shp in enumerate(spec.shapes):
        a = rs.uniform(0, 1, (B,) + tuple(shp))
        if m in binary:
            a = (a > 0.5).astype(np.float64)
        xs.append(_t(a))
    return xs


def make_noise(kind, shape, rs):
    """kind 'laplace': u~U(eps−1,1) (torch Laplace.rsample); 'normal': eps~N(0,1)."""
    if kind == "laplace":
        return _t(rs.uniform(F32_EPS - 1.0, 1.0, shape))
    return _t(rs.standard_normal(shape))


def grad_digest(g: torch.Tensor, keep64=False):
    """Compact, order-sensitive fingerprint of a gradient tensor for golden fixtures (keep64: sampled values stay fp64,
    for the fixtures produced by running the reference in double precision)."""
    f = g.detach().reshape(-1).double()
    n = f.numel()
    idx = torch.linspace(0, n - 1, steps=min(n, 48)).long()
    return {"norm": f.norm().item(), "sum": f.sum().item(), "idx": idx,
            "vals": f[idx].clone() if keep64 else f[idx].float().clone(), "numel": n}
