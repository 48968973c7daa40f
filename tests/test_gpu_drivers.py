"""Training / evaluation drivers end to end on the GPU: config -> model -> epochs through the
CUDA path -> best-validation checkpoint with reference state_dict keys -> K-sample likelihood evaluation."""
import argparse
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(model, **kw):
    c = dict(model=model, obj="dreg", looser=True, K=5, latent_dim=8, dist="laplace", batch_size=64, learning_rate=1e-3,
             epochs=2, seed=1, warmup=0, beta_prior=1, llik_scaling=0.0)
    c.update(kw)
    a = argparse.Namespace()
    a.__dict__.update(c)
    return a


@pytest.mark.parametrize("model", ["mnist_svhn", "circles_squares"])
def test_train_then_eval(tmp_path, model):
    from mmvae_b200 import drivers
    import mmvae_b200
    args = _cfg(model)
    agg = drivers.run_training(args, str(tmp_path), synthetic=700, log=lambda *_: None)
    assert len(agg["train_loss"]) == 2 and all(np.isfinite(agg["train_loss"])) and all(np.isfinite(agg["test_loss"]))
    assert agg["train_loss"][1] < agg["train_loss"][0]                       # it learns
    assert os.path.exists(tmp_path / "model.pt") and os.path.exists(tmp_path / "losses.pt")
    saved = json.load(open(tmp_path / "args.json"))
    assert saved["model"] == model and saved["K"] == 5
    # the checkpoint has the reference's keys: it loads strictly into a freshly built model
    fresh = getattr(mmvae_b200, "VAE_" + model)(args).cuda()
    fresh.load_state_dict(torch.load(tmp_path / "model.pt"), strict=True)
    assert any(k.startswith("vaes.0.encoder") for k in fresh.state_dict()) and "_pz_params.0" in fresh.state_dict()
    res = drivers.run_likelihoods(args, str(tmp_path), K=40, repeats=2, synthetic=700, batch_size=100, batch_size_K=20,
                                  max_batches=2, log=lambda *_: None)
    assert len(res) == 2
    for r in res:
        assert {"likelihood", "cond_likelihood_0_1", "cond_likelihood_1_0"} <= set(r["mean"])
        assert all(np.isfinite(v) for v in r["mean"].values())
    assert res[0]["mean"]["likelihood"] != res[1]["mean"]["likelihood"]      # independent draws per repeat
    assert os.path.exists(tmp_path / "likelihoods_K40.json")


def test_lr_change_drops_and_recaptures_graph():
    from mmvae_b200 import drivers
    from mmvae_b200.train import Trainer
    args = _cfg("circles_squares")
    model = drivers.build_model(args, torch.device("cuda"))
    tr = Trainer(model, "m_dreg_looser", K=4, distributed=False)
    x = [torch.rand(32, 1, 32, 32, device="cuda") for _ in range(2)]
    tr.capture(x, warmup=1)
    assert tr.graph_ready(x) and not tr.graph_ready([t[:7] for t in x])
    tr.set_lr(1e-4)
    assert not tr.graph_ready(x)
    tr.capture(x, warmup=1)
    p0 = tr.fp.flat.clone()
    tr.step_graph(x)
    torch.cuda.synchronize()
    step = (tr.fp.flat - p0).abs().max().item()
    assert 0 < step <= 1.01e-4 * 3                                           # Adam's first-steps magnitude ~ lr


def test_jmvae_nf_driver_follows_the_warmup_schedule(tmp_path, monkeypatch):
    """main.py:155,171 through the driver for a JMVAE-NF run with warmup > 0 (the reference's jnf configs use 15 or 30): while
    epoch < warmup the joint ELBO trains the joint encoder and the decoders; from epoch >= warmup they are frozen
    (objectives.py:180-183) and only the unimodal encoders + flows move; model.step(epoch) is called after every epoch; the
    training and the validation loss of an epoch are the SAME objective (same epoch / warmup / beta_prior), so they stay of one
    size across the phase change."""
    from mmvae_b200 import drivers
    holder, snaps = {}, []
    orig = drivers.build_model

    def build(args, device):
        m = orig(args, device)
        holder["m"] = m
        real_step = m.step
        snap = lambda mod: torch.cat([p.detach().reshape(-1).clone() for p in mod.parameters()])

        def step(epoch):
            snaps.append((epoch, snap(m.joint_encoder), snap(m.vaes[1].decoder), snap(m.vaes[0].encoder)))
            return real_step(epoch)
        m.step = step
        return m
    monkeypatch.setattr(drivers, "build_model", build)
    args = _cfg("jnf_mnist_svhn_dcca", obj="jmvae_nf", looser=False, K=1, latent_dim=20, dist="normal", epochs=4, warmup=3,
                num_hidden_layers=1, beta_kl=1.0, decrease_beta_kl=1.0, fix_jencoder=True, fix_decoders=True, no_recon=True,
                dcca=True, dim_dcca=9, flow="maf", no_nf=False, recon_losses=["normal", "normal"])
    agg = drivers.run_training(args, str(tmp_path), synthetic=400, log=lambda *_: None)
    assert [e for e, *_ in snaps] == [1, 2, 3, 4]                                   # model.step(epoch) after every epoch
    assert all(np.isfinite(agg["train_loss"])) and all(np.isfinite(agg["test_loss"]))
    (_, j1, d1, e1), (_, j2, d2, e2), (_, j3, d3, e3), (_, j4, d4, e4) = snaps
    assert (j2 - j1).abs().max().item() > 0 and (d2 - d1).abs().max().item() > 0    # phase 1: the joint ELBO trains them
    assert torch.equal(j4, j3) and torch.equal(d4, d3)                              # phase 2 (epoch 4 >= warmup ... epoch 3 too)
    assert torch.equal(j3, j2) and torch.equal(d3, d2)                              # epoch 3 = first frozen epoch
    assert (e4 - e3).abs().max().item() > 0                                         # the unimodal encoders keep training
    for tr_l, te_l in zip(agg["train_loss"], agg["test_loss"]):                     # one objective for both numbers
        assert 0.2 < tr_l / te_l < 5.0, (agg["train_loss"], agg["test_loss"])
