"""Multimodal objectives with the reference's names, signatures and return conventions (objectives.py).

    objective(model, x, K, epoch, warmup, beta_prior) -> (scalar to MAXIMISE, dict)        main.py:155-156

Each function accepts the six positionals main.py:155 passes and the keyword form of main.py:185.
The arithmetic runs in the fused kernels of libmmvae_b200 through `model.objective(...)`; a model that is not one of
this package's classes is rejected (there is no eager fallback).

Reference quirks kept: A1 the DReG hook never fires -> weights are constants and q-params are
detached inside log q; A2 `m_iwae*` implement the intended maths with `model.lik_scaling`; A3 `m_iwae_looser`
returns a bare tensor in the reference — here `(tensor, {})` unless strict_reference_returns=True.
"""
import math

import torch

from .models import MMVAE


def compute_microbatch_split(x, K, budget=1e8):
    """objectives.py:10-17: number of datapoints per micro-batch so that K*sum(D_m) stays under `budget` floats
    (the reference sized it "for 12Gb cuda memory"; B200 has 180 GB, so callers may raise the budget)."""
    B = x[0].size(0)
    S = sum(1.0 / (K * math.prod(_x.size()[1:])) for _x in x)
    S = int(budget * S)
    assert S > 0, "Cannot fit individual data in memory, consider smaller K"
    return min(B, S)


def _require_mmvae(model):
    if not isinstance(model, MMVAE):
        raise TypeError(f"mmvae_b200 objectives need an mmvae_b200 MMVAE model, got {type(model).__name__}")


def _split_sum(model, x, K, mode, noise=None):
    """Sum of the objective over micro-batches (objectives.py:394-397,429-433); one chunk in every shipped config."""
    S = compute_microbatch_split(x, K)
    B = x[0].size(0)
    if S >= B:
        return model.objective(x, K, mode, noise)
    total, aux = 0, {}
    for b0 in range(0, B, S):
        nz = None if noise is None else [n[:, b0:b0 + S] for n in noise]
        val, aux = model.objective([_x[b0:b0 + S] for _x in x], K, mode, nz)
        total = total + val
    return total, aux


def m_elbo(model, x, K=1, beta=1000, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:87-111."""
    _require_mmvae(model)
    loss, aux = model.objective(x, K, "elbo", noise)
    terms = aux["terms"]
    M = model.mod
    details = {f"lpx_zs{i // 2}{i % 2}": terms[i] for i in range(min(4, M * M))}      # objectives.py:109-110
    return loss, details


def m_iwae(model, x, K=1, beta=0, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:333-340 (+ :117-131)."""
    _require_mmvae(model)
    loss, _ = _split_sum(model, x, K, "iwae", noise)
    return loss, {}


def m_iwae_looser(model, x, K=1, beta=0, epoch=1, warmup=0, beta_prior=1, noise=None,
                  strict_reference_returns=False):
    """objectives.py:361-369 (+ :343-359)."""
    _require_mmvae(model)
    loss, _ = _split_sum(model, x, K, "iwae_looser", noise)
    return loss if strict_reference_returns else (loss, {})


def m_dreg(model, x, K=1, beta=0, warmup=0, epoch=1, beta_prior=1, noise=None):
    """objectives.py:391-403 (+ :372-389): softmax over all M*K samples."""
    _require_mmvae(model)
    loss, _ = _split_sum(model, x, K, "dreg", noise)
    return loss, {}


def m_dreg_looser(model, x, K=1, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:425-438 (+ :406-422): softmax over K per expert — the shipped MNIST-SVHN training objective."""
    _require_mmvae(model)
    loss, _ = _split_sum(model, x, K, "dreg_looser", noise)
    return loss, {}


def m_jmvae_nf(model, x, K=1, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:179-220: -sum_m s_m * recon_loss_m - beta_reg * (beta_prior * KL(q(z|x,y)||p) + reg), with the
    freezing toggles of :180-183 and the linear warm-up of :185-188 (the wandb.log of :189 is dropped — SURVEY A15).
    noise: None or dict(eps1, eps2, eps_uni) to replay recorded draws."""
    from .jmvae_nf import JMVAE_NF
    if not isinstance(model, JMVAE_NF):
        raise TypeError(f"m_jmvae_nf needs an mmvae_b200 JMVAE_NF model, got {type(model).__name__}")
    return model.objective(x, epoch=epoch, warmup=warmup, beta_prior=beta_prior, noise=noise)


def m_self_built(model, x, K=1, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:481-483: the model computes its own ELBO (MVAE.forward, mvae.py:74-143; MOEPOE.forward, moepoe.py:83-135)."""
    from .mvae import MOEPOE, MVAE
    if not isinstance(model, (MVAE, MOEPOE)):
        raise TypeError(f"m_self_built needs an mmvae_b200 MVAE / MOEPOE model, got {type(model).__name__}")
    return model.forward(x, noise=noise)["elbo"], {}


def m_elbo_nf(model, x, K=1, epoch=1, warmup=0, beta_prior=1, noise=None):
    """objectives.py:463-479: ELBO of the MMVAE-NF with the KL developed (MMVAE_NF.objective)."""
    from .mmvae_nf import MMVAE_NF
    if not isinstance(model, MMVAE_NF):
        raise TypeError(f"m_elbo_nf needs an mmvae_b200 MMVAE_NF model, got {type(model).__name__}")
    return model.objective(x, noise=noise), {}
