"""GPU parity of the JMVAE / JMVAE-NF path: flows, m_jmvae_nf and the
K-sample evaluators against the fp64 oracle and the golden fixtures of the unmodified reference.
Tolerances: loss 1e-5 relative to the fp32 reference and 1e-6 to the fp64 oracle; gradients 5e-5 per tensor against the
fp64 oracle and 1e-4 (+ the reference's own measured fp32 noise) against the reference digests; likelihoods 1e-3 nats."""
import math

import pytest
import torch

from oracle import bivae_oracle as O
from oracle import synth
from tests.helpers import branch_consistent, build_jmvae_nf, leaf_sd, load_golden, rel_err, relu_tap

pytestmark = pytest.mark.gpu
CASES = ["jnf_maf_dcca", "jnf_iaf_recon", "jnf_maf_recon", "jmvae_nonf"]
PKW = {"jnf_maf_dcca": dict(flow="maf", dcca=True), "jnf_iaf_recon": dict(flow="iaf"), "jnf_maf_recon": dict(flow="maf"),
       "jmvae_nonf": dict(no_nf=True)}


def _noise(g):
    return {"eps1": g["eps1"].cuda(), "eps2": g["eps2"].cuda(),
            "eps_uni": None if g["eps_uni"] is None else [e.cuda() for e in g["eps_uni"]]}


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("frozen", [False, True])
def test_m_jmvae_nf_matches_reference_and_f64_oracle(name, frozen):
    from mmvae_b200 import objectives as J
    g, spec, sd = load_golden(name)
    res = g[f"frozen{int(frozen)}"]
    model = build_jmvae_nf(spec, sd, no_recon=g["no_recon"], fix_jencoder=frozen, fix_decoders=frozen, **PKW[name])
    x = [t.cuda() for t in g["x"]]
    with relu_tap(model) as tap:
        loss, det = J.m_jmvae_nf(model, x, 1, 1, 0, 1, noise=_noise(g))        # six positionals, as main.py:155
        loss.backward()
    ref = res["loss"].item()
    for k in ("loss_0", "loss_1", "kld_prior", "reg", "kld_0", "kld_1"):
        a, b = float(det[k]), res["details"][k].item()
        assert abs(a - b) <= 1e-4 * max(abs(b), 1.0), (k, a, b)
    fp = ("joint_encoder.",) + tuple(f"vaes.{m}.decoder." for m in range(spec.M)) if frozen else ()
    dd = lambda t: None if t is None else t.double()

    def run_oracle():
        lsd = leaf_sd(sd, dtype=torch.float64, frozen_prefixes=fp)
        l, _ = O.m_jmvae_nf(spec, lsd, [t.double() for t in g["x"]], dd(g["eps1"]), dd(g["eps2"]),
                            None if g["eps_uni"] is None else [e.double() for e in g["eps_uni"]], epoch=1, warmup=0,
                            no_recon=g["no_recon"])
        l.backward()
        return l.item(), {k: v.grad for k, v in lsd.items()}

    # the fp64 oracle on the implementation's ReLU branches; delta / dl = what units on the other side of a kink change (zero
    # when there are none)
    l64, g64, delta, dl = branch_consistent(run_oracle, tap.masks)
    assert abs(loss.item() - ref) <= 1e-5 * abs(ref) + dl, (loss.item(), ref)
    assert abs(loss.item() - l64) <= 1e-6 * abs(l64)
    grads = {k: p.grad for k, p in model.named_parameters()}
    for k, d in res["grads"].items():
        ours = grads.get(k)
        if d["norm"] == 0.0:
            assert ours is None or ours.norm().item() == 0.0, k
            continue
        ours = ours.detach().reshape(-1).double().cpu()
        exact = g64[k].reshape(-1)
        idx, ref32 = d["idx"], d["vals"].double()
        scale = max(exact[idx].norm().item(), 1e-3 * exact.norm().item())
        e_exact = (ours[idx] - exact[idx]).norm().item()
        assert e_exact <= 5e-5 * scale, f"{name}: {k} off the fp64 oracle by {e_exact / scale:.2e}"
        kink = delta[k].reshape(-1)[idx] if k in delta else torch.zeros_like(ref32)
        ref_noise = (ref32 - (exact[idx] - kink)).norm().item()      # the fp32 reference against the oracle on ITS branches
        e_ref = (ours[idx] - ref32).norm().item()
        assert e_ref <= 1e-4 * scale + 1.05 * (ref_noise + kink.norm().item()), f"{name}: {k} off the reference by {e_ref / scale:.2e}"
    if frozen:                                   # objectives.py:180-183: frozen sub-modules receive no gradient
        assert all(p.grad is None or p.grad.abs().max().item() == 0 for p in model.joint_encoder.parameters())
        assert all(p.grad is None or p.grad.abs().max().item() == 0 for p in model.vaes[1].decoder.parameters())


@pytest.mark.parametrize("name", CASES)
def test_jmvae_nf_evaluators_match_reference(name):
    g, spec, sd = load_golden(name)
    model = build_jmvae_nf(spec, sd, no_recon=g["no_recon"], **PKW[name])
    data = [t.cuda() for t in g["x"]]
    K, bs = g["K"], g["bs"]
    out = model.compute_joint_likelihood(data, K=K, batch_size_K=bs, noise=g["joint"]["noise"].cuda())
    assert abs(out["likelihood"].item() - g["joint"]["likelihood"].item()) < 1e-3
    for c in range(2):
        r = g[f"joint_from_{c}"]
        out = model.compute_joint_ll_from_uni(data, c, K=K, batch_size_K=bs, noise=r["noise"].cuda())
        assert abs(out[f"joint_ll_from_{c}"].item() - r["value"].item()) < 1e-3, c
    met, per = model.compute_conditional_likelihood(data, 0, 1, K=K, batch_size_K=bs, noise=g["cond01"]["noise"].cuda())
    assert torch.allclose(per, g["cond01"]["per"], atol=1e-3, rtol=0)
    assert abs(met["cond_likelihood_0_1"].item() - g["cond01"]["mean"].item()) < 1e-3


def test_forward_tuple_and_step():
    g, spec, sd = load_golden("jnf_maf_recon")
    model = build_jmvae_nf(spec, sd, no_recon=False, flow="maf", decrease_beta_kl=0.5, warmup=2)
    x = [t.cuda() for t in g["x"]]
    qz_xy, recons, z_xy = model(x, noise=g["eps1"].cuda())
    B, L = g["eps1"].shape
    assert z_xy.shape == (B, L) and qz_xy.loc.shape == (B, L)
    assert recons[0].shape == x[0].shape and recons[1].shape == x[1].shape
    mu, std = O.joint_encode(spec, sd, g["x"])
    assert rel_err(qz_xy.loc, mu) < 1e-5 and rel_err(qz_xy.scale, std) < 1e-5
    assert rel_err(z_xy, mu + g["eps1"] * std) < 1e-5
    mu2, std2 = model.joint_encoder(x)                      # joint_encoders.py:76-82: (mu, std)
    assert rel_err(mu2, mu) < 1e-5 and rel_err(std2, std) < 1e-5
    model.step(1)
    assert model.beta_kl == 1.0                             # jmvae_nf.py:189-193: only from the warm-up epoch on
    model.step(2)
    assert model.beta_kl == 0.5


@pytest.mark.parametrize("frozen", [True, False])
def test_trainer_graph_replay_after_eager_steps(frozen):
    """Config C3 (jmvae_nf_dcca.json) through train.Trainer: eager steps followed by CUDA-graph capture and replay.  With the
    joint encoder trainable its parameters are used twice per step (objective + compute_kld); the model must not keep
    the autograd graph of the previous step alive (stored posterior parameters are detached) or the capture fails."""
    import argparse
    import mmvae_b200
    from mmvae_b200.train import Trainer
    params = argparse.Namespace(latent_dim=20, dist="normal", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                num_hidden_layers=1, beta_kl=1.0, decrease_beta_kl=1.0, fix_jencoder=frozen, fix_decoders=frozen,
                                no_recon=True, warmup=0, dcca=True, dim_dcca=9, flow="maf", no_nf=False, device="cuda",
                                data_path="", epochs=1)
    torch.manual_seed(1)
    model = mmvae_b200.VAE_jnf_mnist_svhn_dcca(params).cuda()
    tr = Trainer(model, "m_jmvae_nf", K=1, lr=1e-4, distributed=False)
    g = torch.Generator().manual_seed(3)
    x = [torch.rand(32, 1, 28, 28, generator=g).cuda(), torch.rand(32, 3, 32, 32, generator=g).cuda()]
    first = [tr.step(x).item() for _ in range(3)]
    tr.capture(x)
    later = [tr.step_graph(x).item() for _ in range(20)]
    assert all(map(math.isfinite, first + later))
    n = sum(p.numel() for p in model.joint_encoder.parameters() if p.grad is not None and p.grad.abs().sum() > 0)
    assert (n == 0) == frozen                                    # frozen phase: no gradient reaches the joint encoder
    assert sum(later[-5:]) / 5 > sum(first) / 3                  # the objective is maximised (same batch)


@pytest.mark.parametrize("flow", ["maf", "iaf"])
@pytest.mark.parametrize("blocks", [1, 2, 3])
def test_flows_match_oracle_and_invert(flow, blocks):
    spec = O.spec_jnf_mnist_svhn(20, flow, dcca=False, n_made_blocks=blocks)
    sd = synth.make_state_dict(spec, 7)
    model = build_jmvae_nf(spec, sd, flow=flow)
    fl = model.vaes[0].flow
    R, L = 37, 20
    x = torch.randn(R, L, generator=torch.Generator().manual_seed(1))
    fwd = O.maf_forward if flow == "maf" else O.iaf_forward
    inv = O.maf_inverse if flow == "maf" else O.iaf_inverse
    sd64 = {k: v.double() for k, v in sd.items()}
    xd = x.double().requires_grad_()
    u_ref, l_ref = fwd(spec, sd64, "vaes.0.flow.", xd)
    xg = x.cuda().requires_grad_()
    o = fl(xg)
    assert rel_err(o.out, u_ref) < 2e-6 and rel_err(o.log_abs_det_jac, l_ref) < 2e-6
    w = torch.randn(R, L, generator=torch.Generator().manual_seed(2))
    (gx,) = torch.autograd.grad((o.out * w.cuda()).sum() + o.log_abs_det_jac.sum(), xg)
    (gr,) = torch.autograd.grad((u_ref * w.double()).sum() + l_ref.sum(), xd)
    assert rel_err(gx, gr) < 1e-5
    yd = x.double().requires_grad_()
    z_ref, li_ref = inv(spec, sd64, "vaes.0.flow.", yd)
    yg = x.cuda().requires_grad_()
    oi = fl.inverse(yg)
    assert rel_err(oi.out, z_ref) < 1e-5 and rel_err(oi.log_abs_det_jac, li_ref) < 1e-5
    (gy,) = torch.autograd.grad((oi.out * w.cuda()).sum() + oi.log_abs_det_jac.sum(), yg)
    (gyr,) = torch.autograd.grad((z_ref * w.double()).sum() + li_ref.sum(), yd)
    assert rel_err(gy, gyr) < 1e-4
    # round trip: inverse(forward(x)) = x and the two log-determinants cancel
    with torch.no_grad():
        back = fl.inverse(o.out.detach())
    assert rel_err(back.out, x) < 1e-4
    assert (back.log_abs_det_jac + o.log_abs_det_jac.detach()).abs().max().item() < 1e-3
