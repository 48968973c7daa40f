"""CPU, world_size 2, gloo: the multi-rank host logic of mmvae_b200/parallel.py.
The per-rank COMPUTE is stood in for by the CPU oracle (test infrastructure); what is under test is the sharding, the
slicing of the shared random draws, the gradient all-reduce(SUM) and the (max, sum-exp) merge: N ranks must reproduce
the 1-rank result."""
import math
import os
import socket
import time

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import bivae_oracle as O
from oracle import synth

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _plain(v):
    """tensors -> nested lists (a tensor put on a multiprocessing queue dies with its producer)."""
    if torch.is_tensor(v):
        return ("__t__", v.detach().double().tolist())
    if isinstance(v, (tuple, list)):
        return [_plain(u) for u in v]
    return v


def _tensors(v):
    if isinstance(v, tuple) and len(v) == 2 and v[0] == "__t__":
        return torch.tensor(v[1], dtype=torch.float64).float()
    if isinstance(v, list):
        return [_tensors(u) for u in v]
    return v


def _run(fn, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, port, q) + args) for r in range(WORLD)]
    for p in procs:
        p.start()
    out, t0 = {}, time.time()
    while len(out) < WORLD and time.time() - t0 < 240:          # drain BEFORE joining: a full pipe blocks the producer
        if not q.empty():
            r, v = q.get()
            out[r] = _tensors(v)
        elif any(p.exitcode not in (None, 0) for p in procs):
            break
        else:
            time.sleep(0.05)
    for p in procs:
        p.join(30)
        if p.is_alive():
            p.kill()
    assert all(p.exitcode == 0 for p in procs) and len(out) == WORLD, [p.exitcode for p in procs]
    return out


def _entry(fn, rank, port, q, *args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        q.put((rank, _plain(fn(rank, *args))))
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ workers
def _w_lse(rank):
    from mmvae_b200 import parallel as P
    g = torch.Generator().manual_seed(0)
    a = torch.randn(5, 37, generator=g) * 30 - 3000
    k0, k1 = P.shard_bounds(37, rank, WORLD)
    sh = a[:, k0:k1]
    mx = sh.max(1).values
    se = torch.exp(sh - mx[:, None]).sum(1)
    ll = P.finish_lse(*P.combine_lse(mx, se), 37)
    # one rank with an empty shard
    mx2 = mx if rank == 0 else torch.full((5,), -math.inf)
    se2 = se if rank == 0 else torch.zeros(5)
    ll2 = P.finish_lse(*P.combine_lse(mx2, se2), 37)
    per = torch.arange(*P.shard_bounds(7, rank, WORLD)).float()
    return ll, ll2, P.gather_datapoints(per, 7)


def _w_train(rank):
    """Batch-sharded m_dreg_looser: sum of per-shard gradients (all-reduce SUM of one flat bucket) == full batch."""
    from mmvae_b200 import parallel as P
    spec = O.spec_circles(2, "normal")
    sd = synth.make_state_dict(spec, 3)
    B, K = 6, 4
    x = synth.make_inputs(spec, B, seed=4)
    import numpy as np
    rs = np.random.RandomState(5)
    noise = [synth.make_noise("normal", (K, B, 2), rs) for _ in range(2)]           # the FULL draw, same on all ranks
    b0, b1 = P.shard_bounds(B, rank, WORLD)
    leaves = {k: v.clone().requires_grad_(k.endswith(("weight", "bias"))) for k, v in sd.items()}
    loss = O.m_dreg_looser(spec, leaves, P.shard_batch(x, rank, WORLD), K, [n[:, b0:b1] for n in noise])
    loss.backward()
    keys = sorted(k for k, v in leaves.items() if v.grad is not None)
    flat = torch.cat([leaves[k].grad.reshape(-1) for k in keys])
    P.allreduce_sum_(flat)
    tot = P.allreduce_sum_(loss.detach().clone().reshape(1))
    return flat[::97].clone(), tot


class _OracleModel:
    """Duck-typed stand-in for mmvae_b200.MMVAE whose compute is the CPU oracle (tests only)."""

    def __init__(self, spec, sd):
        self.spec, self.sd, self.mod = spec, sd, spec.M
        self.params = type("P", (), {"latent_dim": spec.latent_dim, "dist": spec.dist})()
        self.vaes = [type("V", (), {"dist": spec.dist})() for _ in range(spec.M)]

    @staticmethod
    def _used(K, bs):
        return (K // bs) * bs

    def _joint_latents(self, data, K, bs, mask, noise):
        mus, stds = zip(*[(lambda m, lv: (m, O.posterior_std(self.spec, lv)))(*O.encode(self.spec, self.sd, m, data[m]))
                          for m in range(2)])
        zs = [O.lat_rsample(self.spec.dist, mus[m].unsqueeze(0), stds[m].unsqueeze(0), noise[m]) for m in range(2)]
        bern = mask.reshape(K, -1, 1)
        z = (bern * zs[0] + (1 - bern) * zs[1]).permute(1, 0, 2)[:, :self._used(K, bs)]
        return z, torch.stack(mus), torch.stack(stds)

    def _a(self, z, data, mods, mu, std, with_prior):
        B, Kc, L = z.shape
        a = 0
        for d in mods:
            r = O.decode(self.spec, self.sd, d, z.reshape(B * Kc, L)).reshape((B, Kc) + tuple(self.spec.shapes[d]))
            a = a + O._lpx_unscaled(self.spec, d, r, data[d].unsqueeze(1))
        if with_prior:
            ploc, pscale = O.pz_params(self.spec, self.sd, z)
            lq = torch.stack([O.lat_log_prob(self.spec.dist, z, mu[e].unsqueeze(1), std[e].unsqueeze(1)).sum(-1)
                              for e in range(mu.shape[0])])
            a = a + O.lat_log_prob(self.spec.dist, z, ploc, pscale).sum(-1) - torch.logsumexp(lq, 0) / 2
        return a

    def _is_loglik(self, z, data, mods, qmode, mu=None, std=None, K_total=None, with_prior=True, partial=False):
        a = self._a(z, data, list(mods), mu, std, with_prior)
        mx = a.max(1).values
        se = torch.exp(a - mx[:, None]).sum(1)
        return (mx, se) if partial else mx + torch.log(se) - math.log(K_total)

    def compute_joint_likelihood(self, data, K, bs, mask=None, noise=None):
        z, mu, std = self._joint_latents(data, K, bs, mask, noise)
        self.last_per_datapoint = self._is_loglik(z, data, range(2), "mixture_half", mu, std, K_total=K)
        return {"likelihood": self.last_per_datapoint.mean()}

    def _cond_latents(self, c, x_c, Ku, noise):
        mu, lv = O.encode(self.spec, self.sd, c, x_c)
        std = O.posterior_std(self.spec, lv)
        return (O.lat_rsample(self.spec.dist, mu.unsqueeze(1), std.unsqueeze(1), noise),)

    def compute_conditional_likelihood(self, data, c, g, K, bs, noise=None):
        z = self._cond_latents(c, data[c], self._used(K, bs), noise)[0]
        per = self._is_loglik(z, data, [g], None, K_total=K, with_prior=False)
        return {}, per


def _w_eval(rank, axis):
    from mmvae_b200 import parallel as P
    spec = O.spec_mnist_svhn(20, "laplace")
    sd = synth.make_state_dict(spec, 6)
    B, K, bs = 3, 12, 4
    data = synth.make_inputs(spec, B, seed=7)
    import numpy as np
    rs = np.random.RandomState(8)
    noise = [synth.make_noise("laplace", (K, B, 20), rs) for _ in range(2)]
    mask = torch.from_numpy(rs.binomial(1, 0.5, (K, B))).float()
    cn = synth.make_noise("laplace", (B, K, 20), rs)
    model = _OracleModel(spec, sd)
    with torch.no_grad():
        out, per = P.joint_likelihood(model, data, K, bs, axis=axis, mask=mask, noise=noise)
        _, cper = P.conditional_likelihood(model, data, 0, 1, K, bs, axis=axis, noise=cn)
    return per, cper


# -------------------------------------------------------------------------------------------------- tests
def test_shard_bounds_cover_the_range():
    from mmvae_b200 import parallel as P
    for n in (0, 1, 7, 128, 1000):
        for w in (1, 2, 3, 8):
            b = [P.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(y - x for x, y in b) - min(y - x for x, y in b) <= 1


def test_lse_merge_and_gather_two_ranks():
    out = _run(_w_lse)
    g = torch.Generator().manual_seed(0)
    a = torch.randn(5, 37, generator=g) * 30 - 3000
    full = torch.logsumexp(a, 1) - math.log(37)
    from mmvae_b200 import parallel as P
    k0, k1 = P.shard_bounds(37, 0, WORLD)
    half = torch.logsumexp(a[:, k0:k1], 1) - math.log(37)
    for r in range(WORLD):
        ll, ll2, gathered = out[r]
        assert torch.allclose(ll, full, atol=1e-4, rtol=0)
        assert torch.allclose(ll2, half, atol=1e-4, rtol=0)
        assert torch.equal(gathered, torch.arange(7).float())


def test_batch_sharded_training_gradients_add_up():
    out = _run(_w_train)
    spec = O.spec_circles(2, "normal")
    sd = synth.make_state_dict(spec, 3)
    import numpy as np
    x = synth.make_inputs(spec, 6, seed=4)
    rs = np.random.RandomState(5)
    noise = [synth.make_noise("normal", (4, 6, 2), rs) for _ in range(2)]
    leaves = {k: v.clone().requires_grad_(k.endswith(("weight", "bias"))) for k, v in sd.items()}
    loss = O.m_dreg_looser(spec, leaves, x, 4, noise)
    loss.backward()
    keys = sorted(k for k, v in leaves.items() if v.grad is not None)
    flat = torch.cat([leaves[k].grad.reshape(-1) for k in keys])
    for r in range(WORLD):
        got, tot = out[r]
        assert ((got - flat[::97]).norm() / flat[::97].norm()).item() < 1e-5
        assert abs(tot.item() - loss.item()) <= 1e-5 * abs(loss.item())


@pytest.mark.parametrize("axis", ["batch", "samples"])
def test_sharded_evaluators_reproduce_one_rank(axis):
    out = _run(_w_eval, axis)
    spec = O.spec_mnist_svhn(20, "laplace")
    sd = synth.make_state_dict(spec, 6)
    import numpy as np
    data = synth.make_inputs(spec, 3, seed=7)
    rs = np.random.RandomState(8)
    noise = [synth.make_noise("laplace", (12, 3, 20), rs) for _ in range(2)]
    mask = torch.from_numpy(rs.binomial(1, 0.5, (12, 3))).float()
    cn = synth.make_noise("laplace", (3, 12, 20), rs)
    with torch.no_grad():
        _, per = O.mmvae_joint_likelihood(spec, sd, data, 12, 4, mask, noise)
        _, cper = O.conditional_likelihood(spec, sd, data, 0, 1, 12, 4, cn)
    for r in range(WORLD):
        assert torch.allclose(out[r][0], per, atol=1e-3, rtol=0), (out[r][0], per)
        assert torch.allclose(out[r][1], cper, atol=1e-3, rtol=0)
