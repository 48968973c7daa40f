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


class relu_tap:
    """with relu_tap(model) as tap: <forward + backward on the CUDA path>;  tap.masks = {parameter name: sign pattern of the
    ReLU layer fed by that weight / bias, rows in call order} (ops.RELU_TAP), for oracle.relu_masks()."""

    def __init__(self, model):
        self.model, self.masks = model, None

    def __enter__(self):
        from mmvae_b200 import ops
        ops.RELU_TAP = {}
        return self

    def __exit__(self, *exc):
        from mmvae_b200 import ops
        torch.cuda.synchronize()
        tap, ops.RELU_TAP = ops.RELU_TAP, None
        names = {p.data_ptr(): k for k, p in self.model.named_parameters()}
        self.masks = {names[ptr]: torch.cat(ms).cpu() for ptr, ms in tap.items() if ptr in names}
        return False


def branch_consistent(run_oracle, masks):
    """run_oracle() -> (loss, {name: grad}) evaluates the fp64 oracle (fresh leaves each call).  Returns
    (loss_b, grads_b, delta): the oracle on the ReLU branches of the implementation under test, and
    delta[name] = grads_b[name] - grads_plain[name] -- exactly zero unless some unit's pre-activation is within the
    implementation's forward error of zero, in which case it is the (O(1) per unit) amount by which ANY two finite-precision
    evaluations may legitimately differ; the comparisons against the reference's fixtures widen by it and by nothing else."""
    O.relu_masks(None)
    l_plain, g_plain = run_oracle()
    O.relu_masks(masks)
    try:
        l_b, g_b = run_oracle()
    finally:
        O.relu_masks(None)
    delta = {k: (g_b[k] - g_plain[k]) for k in g_b if g_b[k] is not None and g_plain.get(k) is not None}
    return l_b, g_b, delta, abs(l_b - l_plain)


def check_digests(grads, digests, rtol, what="", delta=None):
    """grads: name → tensor (or None).  digests: from synth.grad_digest on the reference's .grad.  delta: see
    branch_consistent (slack for units on the other side of a ReLU kink; zero when there are none)."""
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
        dk = delta.get(k) if delta else None
        slack_n = dk.norm().item() if dk is not None else 0.0
        slack_i = dk.reshape(-1).double().cpu()[d["idx"]].norm().item() if dk is not None else 0.0
        assert abs(f.norm().item() - ref_norm) <= rtol * ref_norm + slack_n, f"{what}: {k} norm {f.norm().item()} vs {ref_norm}"
        assert err <= rtol * max(scale, 1e-3 * ref_norm) + 1.05 * slack_i, f"{what}: {k} sampled values off by {err / scale:.3e}"


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
