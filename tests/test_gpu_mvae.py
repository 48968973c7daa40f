"""MVAE on the CUDA path (mmvae_b200/mvae.py) against the fixtures produced by the unmodified reference
(tests/golden/mvae_*.pt) and the fp64 oracle (oracle.bivae_oracle.mvae_*): objective `m_self_built` with gradients,
forward outputs, and the K-sample evaluators.  Tolerances: loss / gradients 1e-4 relative vs the fp32 reference
(measured ~1e-6; this objective has no large-magnitude log-weights, so the reference is not noisy here), 2e-5 vs fp64;
log-likelihoods 1e-3 nats."""
import argparse

import pytest
import torch
import torch.nn as nn

from oracle import bivae_oracle as O
from tests.helpers import branch_consistent, check_digests, leaf_sd, load_golden, rel_err, relu_tap

pytestmark = pytest.mark.gpu


def build_mvae(spec, sd):
    import mmvae_b200
    from mmvae_b200 import networks as N
    from mmvae_b200.vaes import VAEConfig, my_VAE
    params = argparse.Namespace(latent_dim=spec.latent_dim, dist="normal", recon_losses=list(spec.recon_losses),
                                llik_scaling=0.0, device="cuda", data_path="", epochs=1)
    vaes = []
    for m in range(spec.M):
        cfg = VAEConfig(tuple(spec.shapes[m]), spec.latent_dim)
        enc = N.Encoder_VAE_MLP(cfg) if spec.enc[m] == "mlp" else N.Encoder_VAE_SVHN(cfg)
        dec = N.Decoder_AE_MLP(cfg) if spec.dec[m] == "mlp" else N.Decoder_VAE_SVHN(cfg)
        vaes.append(my_VAE(cfg, enc, dec))
    model = mmvae_b200.MVAE(params, nn.ModuleList(vaes))
    model.lik_scaling = tuple(spec.lik_scaling)
    model.load_state_dict(sd, strict=True)                     # the reference's state_dict keys
    return model.cuda()


def _compare_with_reference_and_oracle(model, loss, g, name, run_oracle, masks):
    """loss / gradients vs the fp32 reference fixture (1e-4, widened only by the effect of ReLU units on the other side of a
    kink -- zero when there are none) and vs the fp64 oracle on the same branches (1e-6 / 2e-5)."""
    l64, g64, delta, dl = branch_consistent(run_oracle, masks)
    ref = g["train"]["loss"].item()
    assert abs(loss.item() - ref) <= 1e-4 * abs(ref) + dl, (loss.item(), ref)
    grads = {k: p.grad for k, p in model.named_parameters()}
    check_digests(grads, g["train"]["grads"], 1e-4, name, delta)
    assert abs(loss.item() - l64) <= 1e-6 * abs(l64)
    keys = sorted(k for k, v in g64.items() if v is not None)
    ours = torch.cat([grads[k].detach().double().cpu().reshape(-1) for k in keys])
    theirs = torch.cat([g64[k].reshape(-1) for k in keys])
    assert rel_err(ours, theirs) <= 2e-5


@pytest.mark.parametrize("name", ["mvae_ms", "mvae_msf"])
def test_mvae_objective_matches_reference_and_f64_oracle(name):
    from mmvae_b200 import objectives as J
    g, spec, sd = load_golden(name)
    x = [t.cuda() for t in g["x"]]
    model = build_mvae(spec, sd)
    with relu_tap(model) as tap:
        loss, _ = J.m_self_built(model, x, 1, 1, 0, 1, noise=[n.cuda() for n in g["train"]["noise"]])
        loss.backward()

    def run_oracle():
        lsd = leaf_sd(sd, dtype=torch.float64)
        l = O.mvae_forward(spec, lsd, [t.double() for t in g["x"]], [n.double() for n in g["train"]["noise"]])["elbo"]
        l.backward()
        return l.item(), {k: v.grad for k, v in lsd.items()}

    _compare_with_reference_and_oracle(model, loss, g, name, run_oracle, tap.masks)


@pytest.mark.parametrize("name", ["mvae_ms", "mvae_msf"])
def test_mvae_forward_and_evaluators_match_reference(name):
    g, spec, sd = load_golden(name)
    x = [t.cuda() for t in g["x"]]
    K, bs = g["K"], g["bs"]
    model = build_mvae(spec, sd)
    with torch.no_grad():
        fw = model(x, noise=[n.cuda() for n in g["forward"]["noise"]])
        for k in ("joint_mu", "joint_std", "z_joint"):
            assert torch.allclose(fw[k].cpu(), g["forward"][k], rtol=1e-4, atol=1e-5), k
        assert abs(fw["elbo"].item() - g["forward"]["elbo"].item()) <= 1e-4 * abs(g["forward"]["elbo"].item())
        r = model.compute_joint_likelihood(x, K=K, batch_size_K=bs, noise=g["joint"]["noise"].cuda())
        assert abs(r["likelihood"].item() - g["joint"]["likelihood"].item()) < 1e-3
        _, per = O.mvae_joint_likelihood(spec, {k: v.double() for k, v in sd.items()}, [t.double() for t in g["x"]], K, bs,
                                         g["joint"]["noise"].double())
        assert (model.last_per_datapoint.cpu().double() - per).abs().max().item() < 1e-3
        for c in range(spec.M):
            r = model.compute_joint_ll_from_uni(x, c, K=K, batch_size_K=bs, noise=g[f"joint_from_{c}"]["noise"].cuda())
            assert abs(r[f"joint_ll_from_{c}"].item() - float(g[f"joint_from_{c}"]["value"])) < 1e-3


def test_mvae_registry_and_trainer_step():
    """`VAE_mvae_mnist_svhn` (models/__init__.py:22) built from a config Namespace trains through train.Trainer."""
    import mmvae_b200
    from mmvae_b200.train import Trainer
    params = argparse.Namespace(latent_dim=20, dist="normal", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                device="cuda", data_path="", epochs=1)
    torch.manual_seed(1)
    model = mmvae_b200.VAE_mvae_mnist_svhn(params).cuda()
    assert model.modelName == "mvae_mnist_svhn" and abs(model.lik_scaling[0] - 3072 / 784) < 1e-12
    tr = Trainer(model, "m_self_built", K=1, distributed=False)
    g = torch.Generator().manual_seed(3)
    x = [torch.rand(64, 1, 28, 28, generator=g).cuda(), torch.rand(64, 3, 32, 32, generator=g).cuda()]
    losses = [tr.step(x).item() for _ in range(15)]
    assert losses[-1] > losses[0]                                # the ELBO is maximised


def build_mopoe(spec, sd, beta_kl):
    import mmvae_b200
    from mmvae_b200 import networks as N
    from mmvae_b200.vaes import VAEConfig, my_VAE
    params = argparse.Namespace(latent_dim=spec.latent_dim, dist="normal", recon_losses=list(spec.recon_losses),
                                llik_scaling=0.0, device="cuda", data_path="", epochs=1, beta_kl=beta_kl)
    vaes = []
    for m in range(spec.M):
        cfg = VAEConfig(tuple(spec.shapes[m]), spec.latent_dim)
        enc = N.Encoder_VAE_MLP(cfg) if spec.enc[m] == "mlp" else N.Encoder_VAE_SVHN(cfg)
        dec = N.Decoder_AE_MLP(cfg) if spec.dec[m] == "mlp" else N.Decoder_VAE_SVHN(cfg)
        vaes.append(my_VAE(cfg, enc, dec))
    model = mmvae_b200.MOEPOE(params, nn.ModuleList(vaes))
    model.lik_scaling = tuple(spec.lik_scaling)
    model.load_state_dict(sd, strict=True)
    return model.cuda()


@pytest.mark.parametrize("name", ["mopoe_ms", "mopoe_msf"])
def test_mopoe_matches_reference_and_f64_oracle(name):
    """MoPoE (models/moepoe/moepoe.py:83-214): objective with gradients, mixture components, joint_ll_from_uni."""
    from mmvae_b200 import objectives as J
    g, spec, sd = load_golden(name)
    x = [t.cuda() for t in g["x"]]
    K, bs = g["K"], g["bs"]
    model = build_mopoe(spec, sd, g["beta_kl"])
    with relu_tap(model) as tap:
        loss, _ = J.m_self_built(model, x, 1, 1, 0, 1, noise=g["train"]["noise"].cuda())
        loss.backward()

    def run_oracle():
        lsd = leaf_sd(sd, dtype=torch.float64)
        l = O.mopoe_forward(spec, lsd, [t.double() for t in g["x"]], g["train"]["noise"].double(), g["beta_kl"])["elbo"]
        l.backward()
        return l.item(), {k: v.grad for k, v in lsd.items()}

    _compare_with_reference_and_oracle(model, loss, g, name, run_oracle, tap.masks)
    with torch.no_grad():
        fw = model(x, noise=g["forward"]["noise"].cuda())
        for k in ("mus", "log_vars", "z_joint"):
            assert torch.allclose(fw[k].cpu(), g["forward"][k], rtol=1e-4, atol=1e-5), k
        for c in range(spec.M):
            r = model.compute_joint_ll_from_uni(x, c, K=K, batch_size_K=bs, noise=g[f"joint_from_{c}"]["noise"].cuda())
            assert abs(r[f"joint_ll_from_{c}"].item() - float(g[f"joint_from_{c}"]["value"])) < 1e-3
    with pytest.raises(NotImplementedError):
        model.compute_joint_likelihood(x)
