"""MMVAE-NF (`m_elbo_nf`), product-of-experts subset posteriors, their importance-sampled conditional likelihood and the
Hamiltonian sampler on the CUDA path, against the fixtures produced by the unmodified reference
(tests/golden/mmvae_nf_ms.pt, poe_mvae_msf.pt, poe_jnf_maf.pt) and the fp64 oracle.  Tolerances: loss 1e-4 relative vs the
fp32 reference and 2e-6 vs fp64; gradients 1e-4 / 2e-5; log densities 1e-3 absolute on values of magnitude 10^2..10^3
(fp32 reference noise); log-likelihoods 1e-3 nats; sampler positions 1e-4."""
import argparse

import pytest
import torch
import torch.nn as nn

from oracle import bivae_oracle as O
from tests.helpers import branch_consistent, build_jmvae_nf, check_digests, leaf_sd, load_golden, rel_err, relu_tap
from tests.test_gpu_mvae import build_mvae
from tests.test_oracle_golden import oracle_hmc_start

pytestmark = pytest.mark.gpu


def build_mmvae_nf(spec, sd):
    import mmvae_b200
    from mmvae_b200 import networks as N
    from mmvae_b200.vaes import VAE_IAF_Config, my_VAE_IAF
    params = argparse.Namespace(latent_dim=spec.latent_dim, dist="normal", recon_losses=list(spec.recon_losses),
                                llik_scaling=0.0, device="cuda", data_path="", epochs=1)
    vaes = []
    for m in range(spec.M):
        cfg = VAE_IAF_Config(tuple(spec.shapes[m]), spec.latent_dim)
        cfg.n_made_blocks = spec.n_made_blocks
        enc = N.Encoder_VAE_MLP(cfg) if spec.enc[m] == "mlp" else N.Encoder_VAE_SVHN(cfg)
        dec = N.Decoder_AE_MLP(cfg) if spec.dec[m] == "mlp" else N.Decoder_VAE_SVHN(cfg)
        vaes.append(my_VAE_IAF(cfg, enc, dec))
    model = mmvae_b200.MMVAE_NF(params, nn.ModuleList(vaes))
    model.lik_scaling = tuple(spec.lik_scaling)
    model.load_state_dict(sd, strict=True)
    return model.cuda()


def test_mmvae_nf_objective_matches_reference_and_f64_oracle():
    from mmvae_b200 import objectives as J
    g, spec, sd = load_golden("mmvae_nf_ms")
    x = [t.cuda() for t in g["x"]]
    noise = [n.cuda() for n in g["train"]["noise"]]
    model = build_mmvae_nf(spec, sd)
    with relu_tap(model) as tap:
        loss, det = J.m_elbo_nf(model, x, 1, 1, 0, 1, noise=noise)
        assert det == {}
        loss.backward()

    def run_oracle():
        lsd = leaf_sd(sd, dtype=torch.float64)
        l = O.m_elbo_nf(spec, lsd, [t.double() for t in g["x"]], [n.double() for n in g["train"]["noise"]])
        l.backward()
        return l.item(), {k: v.grad for k, v in lsd.items()}

    l64, g64, delta, dl = branch_consistent(run_oracle, tap.masks)
    ref = g["train"]["loss"].item()
    assert abs(loss.item() - ref) <= 1e-4 * abs(ref) + dl, (loss.item(), ref)
    grads = {k: p.grad for k, p in model.named_parameters()}
    check_digests(grads, g["train"]["grads"], 1e-4, "mmvae_nf_ms", delta)
    assert abs(loss.item() - l64) <= 2e-6 * abs(l64), (loss.item(), l64)
    keys = sorted(k for k, v in g64.items() if v is not None)
    ours = torch.cat([grads[k].detach().double().cpu().reshape(-1) for k in keys])
    theirs = torch.cat([g64[k].reshape(-1) for k in keys])
    assert rel_err(ours, theirs) <= 2e-5
    with torch.no_grad():
        ln_q, zs, recons = model(x, noise=noise)
    for m in range(spec.M):
        assert torch.allclose(ln_q[m].cpu(), g["forward"]["ln_qz_xs"][m], rtol=1e-4, atol=1e-3), m
        assert torch.allclose(zs[m].cpu(), g["forward"]["zs"][m], rtol=1e-4, atol=1e-5), m
        for d in range(spec.M):
            assert recons[m][d].shape == g["x"][d].shape
            ref_s = g["forward"]["recon_sums"][m][d]
            assert abs(recons[m][d].double().sum().item() - ref_s) <= 1e-5 * abs(ref_s)


def test_mmvae_nf_registry_and_trainer_step():
    """`VAE_mmvae_nf_mnist_svhn` (models/__init__.py:21) built from a config Namespace trains through train.Trainer."""
    import mmvae_b200
    from mmvae_b200.train import Trainer
    params = argparse.Namespace(latent_dim=20, dist="normal", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                device="cuda", data_path="", epochs=1)
    torch.manual_seed(1)
    model = mmvae_b200.VAE_mmvae_nf_mnist_svhn(params).cuda()
    assert model.modelName == "mmvae_nf_mnist_svhn" and abs(model.lik_scaling[0] - 3072 / 784) < 1e-12
    assert model.compute_joint_likelihood(None) is None          # mmvae_nf.py:87-88
    tr = Trainer(model, "m_elbo_nf", K=1, distributed=False)
    g = torch.Generator().manual_seed(3)
    x = [torch.rand(64, 1, 28, 28, generator=g).cuda(), torch.rand(64, 3, 32, 32, generator=g).cuda()]
    losses = [tr.step(x).item() for _ in range(15)]
    assert losses[-1] > losses[0]                                # the ELBO is maximised


def _build(g, spec, sd):
    return build_mvae(spec, sd) if g["model"] == "mvae" else build_jmvae_nf(spec, sd, no_recon=True, flow="maf")


@pytest.mark.parametrize("name", ["poe_mvae_msf", "poe_jnf_maf"])
def test_poe_posterior_and_conditional_likelihood(name):
    from mmvae_b200.poe_subset import compute_cond_ll_from_poe_subset
    g, spec, sd = load_golden(name)
    x = [t.cuda() for t in g["x"]]
    K, bs, kind = g["K"], g["bs"], g["model"]
    model = _build(g, spec, sd)
    sd64 = {k: v.double() for k, v in sd.items()}
    x64 = [t.double() for t in g["x"]]
    zq = g["posterior"]["z"].cuda()
    for c in g["posterior"]["cases"]:
        if kind == "mvae":
            with torch.no_grad():
                v = model.compute_poe_posterior(c["subset"], zq, x)
            assert torch.allclose(v.cpu(), c["value"], rtol=1e-5, atol=1e-3)
        else:
            v, gr = model.compute_poe_posterior(c["subset"], zq, x, divide_prior=True, grad=True)
            assert torch.allclose(v.cpu(), c["value"], rtol=1e-5, atol=1e-3)
            assert rel_err(gr, c["grad"]) <= 1e-4
            v64, g64 = O.jmvae_nf_poe_posterior_grad(spec, sd64, c["subset"], g["posterior"]["z"].double(), x64, True)
            assert (v.cpu().double() - v64).abs().max().item() <= 2e-6 * v64.abs().max().item()
            assert rel_err(gr, g64) <= 2e-5
            ex = model.poe_experts(c["subset"], x)                           # cached experts: same numbers
            v_c, g_c = model.compute_poe_posterior(c["subset"], zq, x, True, True, experts=ex)
            assert torch.equal(v_c, v) and torch.equal(g_c, gr)
            with torch.no_grad():
                v2 = model.compute_poe_posterior(c["subset"], zq, x, divide_prior=False, grad=False)
            assert torch.allclose(v2.cpu(), c["value_noprior"], rtol=1e-5, atol=1e-3)
    for c in g["cond"]:
        r = compute_cond_ll_from_poe_subset(model, x, c["subset"], c["gen_mod"], K=K, batch_size_K=bs, i_s=c["i_s"].numpy(),
                                            noise=c["noise"].cuda())
        assert abs(r.item() - float(c["value"])) < 1e-3, (r.item(), float(c["value"]))
        with torch.no_grad():
            _, per = O.cond_ll_from_poe_subset(kind, spec, sd64, x64, c["subset"], c["gen_mod"], K, bs, c["i_s"], c["noise"].double())
        assert (model.last_per_datapoint.cpu().double() - per).abs().max().item() < 1e-3
    with pytest.raises(ValueError):
        compute_cond_ll_from_poe_subset(model, x, g["cond"][0]["subset"], g["cond"][0]["gen_mod"], K=3, batch_size_K=200)


def test_mvae_trimodal_conditional_likelihoods_include_poe_subsets():
    """msf.py:101-108: compute_conditional_likelihoods of the trimodal MVAE adds cond_poe_ll_{0,1,2}; chunked evaluation
    (eval_max_rows) gives the same numbers as one pass."""
    import mmvae_b200
    from mmvae_b200.poe_subset import compute_cond_ll_from_poe_subset
    params = argparse.Namespace(latent_dim=30, dist="normal", recon_losses=["normal"] * 3, llik_scaling=0.0, device="cuda",
                                data_path="", epochs=1)
    torch.manual_seed(5)
    model = mmvae_b200.VAE_mvae_msf(params).cuda()
    g = torch.Generator().manual_seed(7)
    x = [torch.rand(6, 1, 28, 28, generator=g).cuda(), torch.rand(6, 3, 32, 32, generator=g).cuda(),
         torch.rand(6, 1, 28, 28, generator=g).cuda()]
    d = model.compute_conditional_likelihoods(x, K=400, batch_size_K=100)
    for gm in range(3):
        assert torch.isfinite(d[f"cond_poe_ll_{gm}"])
    assert "cond_likelihood_0_1" in d and "cond_lw_subset_0" in d
    noise = torch.randn(6, 400, 30, generator=g).cuda()
    i_s = [0, 1, 1, 0, 1, 0]
    a = compute_cond_ll_from_poe_subset(model, x, [0, 2], 1, K=400, batch_size_K=200, i_s=i_s, noise=noise)
    per_a = model.last_per_datapoint.clone()
    model.eval_max_rows = 800                                    # two datapoints per chunk
    b = compute_cond_ll_from_poe_subset(model, x, [0, 2], 1, K=400, batch_size_K=200, i_s=i_s, noise=noise)
    assert torch.allclose(per_a, model.last_per_datapoint, rtol=0, atol=2e-3) and abs(a.item() - b.item()) < 1e-3


@pytest.mark.parametrize("graph", [False, True])
def test_hmc_poe_sampler_matches_reference(graph):
    g, spec, sd = load_golden("poe_jnf_maf")
    h = g["hmc"]
    x = [t.cuda() for t in g["x"]]
    model = _build(g, spec, sd)
    z = model.sample_from_poe_subset(h["subset"], x, mcmc_steps=h["steps"], n_lf=h["n_lf"], eps_lf=h["eps_lf"], K=h["K"],
                                     divide_prior=True, indices=h["indices"].numpy(), noise=[n.cuda() for n in h["init_noise"]],
                                     gammas=h["gammas"], accepts=h["accepts"], graph=graph)
    assert z.shape == h["z"].shape
    assert torch.allclose(z.cpu(), h["z"], rtol=1e-4, atol=1e-4), (z.cpu() - h["z"]).abs().max()
    sd64 = {k: v.double() for k, v in sd.items()}
    x64 = [t.double() for t in g["x"]]
    z_init = oracle_hmc_start(spec, sd64, x64, h)
    z64, acc = O.jmvae_nf_hmc_poe_subset(spec, sd64, h["subset"], x64, h["K"], z_init, [t.double() for t in h["gammas"]],
                                         [t.double() for t in h["accepts"]], h["n_lf"], h["eps_lf"], True)
    assert (z.cpu().double() - z64).abs().max().item() < 1e-4
    assert torch.equal(model.last_acceptance.cpu().reshape(-1).double() * h["steps"], acc)


def test_hmc_graph_replay_equals_eager_and_moves_towards_the_posterior():
    """Same draws through the CUDA-graph leapfrog and the eager one: identical chains; and the sampler raises the PoE
    log-density of the chains it started from the mixture of experts (default reference settings, shortened)."""
    g, spec, sd = load_golden("poe_jnf_maf")
    model = _build(g, spec, sd)
    gen = torch.Generator().manual_seed(11)
    x = [torch.rand(16, 1, 28, 28, generator=gen).cuda(), torch.rand(16, 3, 32, 32, generator=gen).cuda()]
    K, steps, n = 4, 12, 64
    idx = torch.randint(0, 2, (n,), generator=gen).numpy()
    init = [torch.randn(int((idx == m).sum()), 20, generator=gen).cuda() for m in (0, 1)]
    gam = [torch.randn(n, 20, generator=gen) for _ in range(steps)]
    acc = [torch.rand(n, generator=gen) for _ in range(steps)]
    kw = dict(mcmc_steps=steps, n_lf=10, eps_lf=0.01, K=K, divide_prior=True, indices=idx, noise=init, gammas=gam, accepts=acc)
    z_e = model.sample_from_poe_subset([0, 1], x, graph=False, **kw)
    z_g = model.sample_from_poe_subset([0, 1], x, graph=True, **kw)
    assert torch.equal(z_e, z_g)
    rep = [torch.cat([d] * K) for d in x]
    z0 = model.sample_from_moe_subset([0, 1], rep, idx, init)
    with torch.no_grad():
        before = model.compute_poe_posterior([0, 1], z0, rep, True, grad=False)
        after = model.compute_poe_posterior([0, 1], z_g.reshape(n, 20), rep, True, grad=False)
    assert torch.isfinite(after).all() and float(model.last_acceptance.mean()) > 0.5
