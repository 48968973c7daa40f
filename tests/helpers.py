"""Shared test helpers: golden fixture loading and oracle evaluation (test infrastructure)."""
import os

import torch

from oracle import bivae_oracle as O
from oracle import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    g = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    kw = dict(g["spec"])
    kw["shapes"] = [tuple(s) for s in kw["shapes"]]
    spec = O.Spec(**kw)
    sd = synth.make_state_dict(spec, g["wseed"])
    return g, spec, sd


def leaf_sd(sd, dtype=torch.float32, frozen_prefixes=()):
    """Copy of the state dict whose float weights are autograd leaves (masks/prior/first-stage stay constant)."""
    out = {}
    for k, v in sd.items():
        v = v.detach().clone().to(dtype)
        trainable = (k.endswith("weight") or k.endswith("bias")) and "first_encoder" not in k \
            and not any(k.startswith(p) for p in frozen_prefixes)
        out[k] = v.requires_grad_(trainable)
    return out


def check_digests(grads, digests, rtol, what=""):
    """grads: name → tensor (or None).  digests: from synth.grad_digest on the reference's .grad."""
    for k, d in digests.items():
        g = grads.get(k)
        assert g is not None, f"{what}: no gradient for {k}"
        f = g.detach().reshape(-1).double().cpu()
        assert f.numel() == d["numel"], k
        ref_norm = d["norm"]
        err = (f[d["idx"]] - d["vals"].double()).norm().item()
        scale = max(d["vals"].double().norm().item(), 1e-30)
        if ref_norm == 0.0:
            assert f.norm().item() == 0.0, f"{what}: {k} should have zero grad"
            continue
        assert abs(f.norm().item() - ref_norm) <= rtol * ref_norm, f"{what}: {k} norm {f.norm().item()} vs {ref_norm}"
        assert err <= rtol * max(scale, 1e-3 * ref_norm), f"{what}: {k} sampled values off by {err / scale:.3e}"


# ------------------------------------------------------------------------------------------------ GPU helpers
def build_mmvae(spec, sd, device="cuda"):
    """mmvae_b200 MMVAE with the architecture of `spec` and the weights of the reference-keyed state dict `sd`."""
    import argparse
    import torch.nn as nn
    import mmvae_b200
    from mmvae_b200 import networks as N
    from mmvae_b200.vaes import VAEConfig, laplace_VAE, my_VAE
    params = argparse.Namespace(latent_dim=spec.latent_dim, dist=spec.dist, recon_losses=list(spec.recon_losses),
                                llik_scaling=0.0, device=device, data_path="", epochs=1)
    vae_cls = my_VAE if spec.vae == "my_VAE" else laplace_VAE
    vaes = []
    for m in range(spec.M):
        cfg = VAEConfig(tuple(spec.shapes[m]), spec.latent_dim)
        enc = N.Encoder_VAE_MLP(cfg) if spec.enc[m] == "mlp" else N.Encoder_VAE_SVHN(cfg)
        dec = N.Decoder_AE_MLP(cfg) if spec.dec[m] == "mlp" else N.Decoder_VAE_SVHN(cfg)
        vaes.append(vae_cls(cfg, enc, dec))
    model = mmvae_b200.MMVAE(params, nn.ModuleList(vaes))
    model.lik_scaling = tuple(spec.lik_scaling)
    model.load_state_dict(sd, strict=True)
    return model.to(device)


def rel_err(a, b):
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def build_jmvae_nf(spec, sd, device="cuda", **pkw):
    """mmvae_b200 JMVAE_NF with the architecture of `spec` (oracle.spec_jnf_mnist_svhn) and the weights of `sd`."""
    import argparse
    import torch.nn as nn
    from mmvae_b200 import jmvae_nf as J
    from mmvae_b200 import networks as N
    from mmvae_b200 import vaes as V
    params = argparse.Namespace(latent_dim=spec.latent_dim, dist="normal", recon_losses=list(spec.recon_losses),
                                num_hidden_layers=spec.num_hidden_layers, llik_scaling=0.0, device=device, data_path="",
                                epochs=1, beta_kl=1.0, decrease_beta_kl=1.0, fix_jencoder=False, fix_decoders=False,
                                no_recon=True, warmup=0, dcca=False)
    for k, v in pkw.items():
        setattr(params, k, v)
    cfg_cls, vae_cls = {"my_VAE": (V.VAEConfig, V.my_VAE), "my_VAE_IAF": (V.VAE_IAF_Config, V.my_VAE_IAF),
                        "my_VAE_MAF": (V.VAE_MAF_Config, V.my_VAE_MAF)}[spec.vae]
    enc_of = {"mlp": N.Encoder_VAE_MLP, "svhn": N.Encoder_VAE_SVHN}
    pre = [V.VAEConfig(tuple(s), spec.joint_head_latent) for s in spec.shapes]
    joint = N.DoubleHeadJoint(512, pre[0], pre[1], enc_of[spec.joint_heads[0]], enc_of[spec.joint_heads[1]], params)
    vaes, dcca = [], []
    for m in range(spec.M):
        cfg = cfg_cls(tuple(spec.shapes[m]), spec.latent_dim)
        if spec.vae != "my_VAE":
            cfg.n_made_blocks = spec.n_made_blocks
        if spec.enc[m] == "twostep":
            fcfg = V.VAEConfig(tuple(spec.shapes[m]), spec.dcca_latent)
            first = N.LinearCCAEncoder(enc_of[spec.first_enc[m]](fcfg), spec.dcca_latent, spec.dim_dcca)
            dcca.append(first)
            enc = N.TwoStepsEncoder(first, params)
        else:
            enc = enc_of[spec.enc[m]](cfg)
        dec = N.Decoder_AE_MLP(cfg) if spec.dec[m] == "mlp" else N.Decoder_VAE_SVHN(cfg)
        vaes.append(vae_cls(model_config=cfg, encoder=enc, decoder=dec))
    model = J.JMVAE_NF(params, joint, nn.ModuleList(vaes))
    if dcca:
        model.dcca = dcca
    model.lik_scaling = tuple(spec.lik_scaling)
    model.load_state_dict(sd, strict=True)
    return model.to(device)
