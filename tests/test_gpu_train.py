"""Trainer (train.py) on the GPU: the flat-buffer step with gradients sunk straight into the flat gradient buffer and the
per-modality chains on side streams must give the gradients of the plain autograd path, which are pinned to the oracle
(reference main.py:147-160 semantics: loss = -objective, backward, Adam(amsgrad))."""
import copy

import numpy as np
import pytest
import torch

from oracle import bivae_oracle as O
from oracle import synth
from tests.helpers import build_mmvae, rel_err

pytestmark = pytest.mark.gpu


def _setup(B=16, K=6, L=20, seed=3):
    spec = O.spec_mnist_svhn(L, "laplace")
    sd = synth.make_state_dict(spec, seed)
    x = [t.cuda() for t in synth.make_inputs(spec, B, seed=seed + 1)]
    rs = np.random.RandomState(seed + 2)
    noises = [[synth.make_noise("laplace", (K, B, L), rs).cuda() for _ in range(2)] for _ in range(3)]
    return spec, sd, x, noises, K


def _run(spec, sd, x, noises, K, branch_streams, grad_sink, objective="m_dreg_looser", wgrad_stream=True):
    from mmvae_b200.train import Trainer
    model = build_mmvae(spec, sd)
    tr = Trainer(model, objective, K=K, lr=1e-3, amsgrad=True, branch_streams=branch_streams, grad_sink=grad_sink,
                 wgrad_stream=wgrad_stream, distributed=False)
    losses, g0 = [], None
    for nz in noises:
        losses.append(tr.step(x, noise=nz).item())
        if g0 is None:
            torch.cuda.synchronize()
            g0 = tr.fp.grad.clone()
    torch.cuda.synchronize()
    return losses, g0, tr.fp.flat.clone(), tr


@pytest.mark.parametrize("objective", ["m_dreg_looser", "m_iwae", "m_elbo"])
def test_sink_and_streams_match_plain_autograd(objective):
    spec, sd, x, noises, K = _setup()
    l0, g0, p0, _ = _run(spec, sd, x, noises, K, False, False, objective)
    for bs, gs, ws in ((False, True, False), (True, False, False), (True, True, False), (False, True, True),
                       (True, True, True)):
        l1, g1, p1, _ = _run(spec, sd, x, noises, K, bs, gs, objective, ws)
        assert l1 == l0, (bs, gs, ws, l1, l0)
        assert torch.equal(g1, g0), (bs, gs, ws, rel_err(g1, g0))      # same kernels, same order of summation
        assert torch.equal(p1, p0), (bs, gs, ws)


def test_trainer_gradient_matches_oracle():
    spec, sd, x, noises, K = _setup()
    _, g, _, tr = _run(spec, sd, x, noises[:1], K, True, True)
    # fp64 oracle: the fp32 reference arithmetic is itself 1e-4..6e-4 away from fp64 on these gradients
    lsd = {k: v.double().clone().requires_grad_(k.endswith(("weight", "bias"))) for k, v in sd.items()}
    ref = O.m_dreg_looser(spec, lsd, [t.cpu().double() for t in x], K, [n.cpu().double() for n in noises[0]])
    (-ref).backward()
    names = [n for n, p in tr.model.named_parameters() if p.requires_grad]
    flat_ref = torch.cat([lsd[n].grad.reshape(-1) for n in names])
    assert flat_ref.numel() == g.numel()
    assert rel_err(g, flat_ref) < 1e-4           # north_star: gradients within 1e-4 relative in fp32


def test_graph_replay_with_streams_trains():
    from mmvae_b200.train import Trainer
    spec, sd, x, _, K = _setup(B=32, K=8)
    model = build_mmvae(spec, sd)
    tr = Trainer(model, "m_dreg_looser", K=K, distributed=False)
    p_before = tr.fp.flat.clone()
    tr.capture(x, warmup=2)
    losses = [tr.step_graph(x).item() for _ in range(20)]
    torch.cuda.synchronize()
    assert all(np.isfinite(losses))
    assert not torch.equal(tr.fp.flat, p_before)
    assert np.mean(losses[-5:]) > np.mean(losses[:5])                  # the objective is maximised (main.py:156)
    # the same model stepped eagerly afterwards agrees with a graph replay on the same state within RNG noise
    e = tr.step(x).item()
    assert abs(e - losses[-1]) < 0.2 * abs(losses[-1])


def test_streams_full_size_bitwise_c2():
    """C2 shapes (B=128, K=30): with kernels long enough to really overlap, the three-stream schedule must still give
    bit-identical gradients and parameters (a missing dependency would show up as a race here)."""
    spec, sd, x, noises, K = _setup(B=128, K=30, seed=11)
    l0, g0, p0, _ = _run(spec, sd, x, noises, K, False, False)
    for rep in range(2):
        l1, g1, p1, _ = _run(spec, sd, x, noises, K, True, True, wgrad_stream=True)
        assert l1 == l0
        assert torch.equal(g1, g0), rel_err(g1, g0)
        assert torch.equal(p1, p0)


@pytest.mark.parametrize("family,objective,K,dist_", [("VAE_mnist_svhn", "m_dreg_looser", 5, "laplace"),
                                                     ("VAE_mnist_svhn", "m_elbo", 3, "normal"),
                                                     ("VAE_mvae_mnist_svhn", "m_self_built", 1, "normal"),
                                                     ("VAE_moepoe_mnist_svhn", "m_self_built", 1, "normal"),
                                                     ("VAE_mmvae_nf_mnist_svhn", "m_elbo_nf", 1, "normal")])
def test_every_model_family_captures_after_eager_steps(family, objective, K, dist_):
    """drivers.run_training mixes eager steps (ragged batches) with CUDA-graph replay: no model may keep the autograd
    graph of an eager step alive across steps, and the replayed step must keep optimising the same batch."""
    import argparse
    import math
    import mmvae_b200
    from mmvae_b200.train import Trainer
    params = argparse.Namespace(latent_dim=20, dist=dist_, recon_losses=["normal", "normal"], llik_scaling=0.0, device="cuda",
                                data_path="", epochs=1, beta_kl=1.0)
    torch.manual_seed(2)
    model = getattr(mmvae_b200, family)(params).cuda()
    tr = Trainer(model, objective, K=K, lr=1e-4, distributed=False)
    g = torch.Generator().manual_seed(4)
    x = [torch.rand(24, 1, 28, 28, generator=g).cuda(), torch.rand(24, 3, 32, 32, generator=g).cuda()]
    first = [tr.step(x).item() for _ in range(3)]
    tr.capture(x)
    later = [tr.step_graph(x).item() for _ in range(30)]
    assert all(map(math.isfinite, first + later))
    if objective != "m_elbo":                                    # m_elbo's exp(log q_r - log q_d) weights are heavy-tailed at K=3
        assert sum(later[-5:]) / 5 > sum(first) / 3


def test_trainer_passes_the_schedule_and_freezes_like_the_reference():
    """main.py:155,171: objective(model, data, K, epoch, warmup, beta_prior) and model.step(epoch).  m_jmvae_nf trains the joint
    ELBO while epoch < warmup; from epoch >= warmup the joint encoder and the decoders are frozen (objectives.py:180-183) and
    torch's Adam skips them (grad None): they must not move, not even along their momentum; a captured graph is dropped when
    the phase changes; capturing does not advance the optimiser."""
    import numpy as np
    from oracle import bivae_oracle as O
    from oracle import synth
    from tests.helpers import build_jmvae_nf
    from mmvae_b200.train import Trainer
    spec = O.spec_jnf_mnist_svhn(20, "maf", dcca=False)
    sd = synth.make_state_dict(spec, 5)
    model = build_jmvae_nf(spec, sd, flow="maf", fix_jencoder=True, fix_decoders=True, warmup=2)
    tr = Trainer(model, "m_jmvae_nf", K=1, distributed=False)
    x = [t.cuda() for t in synth.make_inputs(spec, 32, seed=6)]
    snap = lambda mod: torch.cat([p.detach().reshape(-1).clone() for p in mod.parameters()])
    # phase 1 (epoch 1 < warmup 2): everything trains, eagerly and through a captured graph
    tr.set_schedule(1, warmup=2, beta_prior=1)
    j0, d0, e0 = snap(model.joint_encoder), snap(model.vaes[1].decoder), snap(model.vaes[0].encoder)
    tr.step(x)
    j1 = snap(model.joint_encoder)
    assert (j1 - j0).abs().max().item() > 0 and (snap(model.vaes[1].decoder) - d0).abs().max().item() > 0
    before, steps_before = tr.fp.flat.clone(), tr.step_dev.item()
    tr.capture(x, warmup=2)
    torch.cuda.synchronize()
    assert torch.equal(tr.fp.flat, before) and tr.step_dev.item() == steps_before       # capture is invisible
    assert tr.graph_ready(x)
    tr.step_graph(x)
    torch.cuda.synchronize()
    assert (snap(model.joint_encoder) - j1).abs().max().item() > 0
    # phase 2 (epoch 2 >= warmup): the graph of phase 1 is gone; joint encoder and decoders stay bit-identical
    tr.set_schedule(2, warmup=2, beta_prior=1)
    assert not tr.graph_ready(x)
    j2, d2, e2 = snap(model.joint_encoder), snap(model.vaes[1].decoder), snap(model.vaes[0].encoder)
    for _ in range(2):
        tr.step(x)
    tr.capture(x, warmup=1)
    tr.step_graph(x)
    torch.cuda.synchronize()
    assert torch.equal(snap(model.joint_encoder), j2) and torch.equal(snap(model.vaes[1].decoder), d2)
    assert (snap(model.vaes[0].encoder) - e2).abs().max().item() > 0                    # the unimodal encoders train
    # the model's own per-epoch schedule reaches the next capture
    model.decrease_beta_kl = 0.5
    model.step(2)
    tr.set_schedule(3, warmup=2, beta_prior=1)
    assert not tr.graph_ready(x)


def test_two_rank_nccl_step_equals_one_rank():
    """The N-rank step on the CUDA path with the NCCL all-reduce (eager and captured in the step graph) against the 1-rank
    step on the concatenated batch (tools/nccl_equality.py).  Needs two GPUs on the box."""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(root, "tools", "nccl_equality.py")], capture_output=True, text=True,
                       timeout=600, cwd=root)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_sink_with_parameters_used_twice_in_one_pass():
    """m_jmvae_nf with a trainable joint encoder and the reconstruction terms on (fix_jencoder = fix_decoders = False,
    no_recon = False): the joint encoder runs in objective() AND compute_kld(), every unimodal encoder in encode_raw() and
    vae(x), every MADE once per flow call.  The first gradient of such a parameter is written by its kernel on the
    weight-gradient stream into p.grad, the second comes back through autograd's `+=` on another stream: the two must be
    ordered (ops._sink).  Gradients and parameters after three steps against plain autograd accumulation on one stream: equal
    to rounding (a parameter with three uses is summed as g1 + (g2 + g3) instead of (g1 + g2) + g3; a LOST contribution would be
    an O(1) error), and identical from run to run; then the same through graph replay."""
    from tests.helpers import build_jmvae_nf
    from mmvae_b200.train import Trainer
    spec = O.spec_jnf_mnist_svhn(20, "maf", dcca=False)
    sd = synth.make_state_dict(spec, 21)
    B = 64
    x = [t.cuda() for t in synth.make_inputs(spec, B, seed=22)]
    rs = np.random.RandomState(23)
    nz = lambda: synth.make_noise("normal", (B, 20), rs).cuda()
    noises = [{"eps1": nz(), "eps2": nz(), "eps_uni": [nz(), nz()]} for _ in range(3)]

    def run(sink, streams):
        model = build_jmvae_nf(spec, sd, flow="maf", fix_jencoder=False, fix_decoders=False, no_recon=False, warmup=0)
        tr = Trainer(model, "m_jmvae_nf", K=1, lr=1e-3, branch_streams=streams, grad_sink=sink, wgrad_stream=streams,
                     distributed=False)
        losses, g0 = [], None
        for nz in noises:
            losses.append(tr.step(x, noise=nz).item())
            if g0 is None:
                torch.cuda.synchronize()
                g0 = tr.fp.grad.clone()
        torch.cuda.synchronize()
        return losses, g0, tr.fp.flat.clone(), tr
    l0, g0, p0, _ = run(False, False)
    assert g0.abs().max().item() > 0
    prev = None
    for rep in range(3):
        l1, g1, p1, tr = run(True, True)
        assert l1[0] == l0[0] and max(abs(a - b) / abs(b) for a, b in zip(l1, l0)) < 1e-6
        assert rel_err(g1, g0) < 1e-7, rel_err(g1, g0)
        assert rel_err(p1, p0) < 1e-6, rel_err(p1, p0)
        assert prev is None or (torch.equal(g1, prev[0]) and torch.equal(p1, prev[1]))       # no race: run-to-run identical
        prev = (g1, p1)
    # graph replay of the same schedule: two replays on the same state as two eager steps of the plain trainer
    _, _, _, plain = run(False, False)
    tr.capture(x, warmup=1)
    for _ in range(2):
        tr.step_graph(x)
        plain.step(x)
    torch.cuda.synchronize()
    assert np.isfinite(float(tr._static_loss))
    # the RNG streams of the two trainers differ (graph-captured noise), so compare the update sizes, not the bits
    assert rel_err(tr.fp.flat, plain.fp.flat) < 5e-2
