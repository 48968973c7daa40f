"""Multi-GPU host logic of the hot path: one process per GPU, torch.distributed (NCCL on the GPU box,
gloo in the CPU tests) for the plumbing.

The path shards without any data-path exchange until the very end:
  * training: every objective is a SUM over datapoints (objectives.py:403,438) -> shard the batch, one all-reduce(SUM)
    of the flat fp32 gradient bucket (train.Trainer), nothing else;
  * evaluation, batch axis: datapoints are independent (mmvae.py:139) -> shard the loader, all-gather the per-datapoint
    ln p_hat (or all-reduce their sum);
  * evaluation, sample axis: ln p_hat = logsumexp_k a_k - ln K is associative -> each rank reduces its K/N importance
    samples to a per-datapoint (max, sum exp) pair, merged with all-reduce(MAX) + all-reduce(SUM): the reference's own
    two-level form (logsumexp over chunks of batch_size_K, then over the chunk results: mmvae.py:168-175).
Random draws (noise, mixture mask) are generated for the FULL problem from one seed and sliced per rank, so an N-rank
run reproduces the 1-rank numbers up to fp32 reassociation.

Everything here works on whatever device the tensors live on; the compute it shards is the CUDA path only.
"""
import math

import torch
import torch.distributed as dist


def rank_world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_bounds(n, rank, world):
    """Contiguous balanced split of range(n): the first n % world shards get one extra element."""
    q, r = divmod(n, world)
    b0 = rank * q + min(rank, r)
    return b0, b0 + q + (1 if rank < r else 0)


def shard_batch(x, rank, world):
    """x: list of [B,...] tensors -> this rank's datapoints."""
    b0, b1 = shard_bounds(x[0].shape[0], rank, world)
    return [t[b0:b1] for t in x]


def allreduce_sum_(t, group=None):
    """In-place SUM all-reduce (a no-op in a single process)."""
    if rank_world(group)[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def combine_lse(mx, se, group=None):
    """Merge per-rank (max, sum exp(a - max)) pairs [B] into the global pair.  Exact in the sense that the result is the
    (max, sum) pair of the union of the shards; -inf maxima (an empty shard) are handled."""
    if rank_world(group)[1] == 1:
        return mx, se
    g = mx.clone()
    dist.all_reduce(g, op=dist.ReduceOp.MAX, group=group)
    scaled = torch.where(torch.isfinite(mx), se * torch.exp(mx - g), torch.zeros_like(se))
    dist.all_reduce(scaled, op=dist.ReduceOp.SUM, group=group)
    return g, scaled


def finish_lse(mx, se, K_total):
    """ln p_hat = max + ln(sum) - ln K."""
    return mx + torch.log(se) - math.log(K_total)


def gather_datapoints(per, B, group=None):
    """per: this rank's per-datapoint values (shard_bounds order) -> all B values on every rank."""
    rank, world = rank_world(group)
    if world == 1:
        return per
    sizes = [shard_bounds(B, r, world) for r in range(world)]
    pad = max(b1 - b0 for b0, b1 in sizes)
    buf = torch.zeros(pad, device=per.device, dtype=per.dtype)
    buf[:per.numel()] = per
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:b1 - b0] for o, (b0, b1) in zip(out, sizes)])


# --------------------------------------------------------------------------------------- sharded evaluators
@torch.no_grad()
def joint_likelihood(model, data, K=1000, batch_size_K=100, axis="batch", mask=None, noise=None, group=None, seed=0):
    """MMVAE.compute_joint_likelihood (mmvae.py:121-177) over all ranks of `group`.
    axis='batch': each rank takes a slice of the datapoints; axis='samples': each rank takes a slice of the K
    importance samples of every datapoint.  mask (K,B) / noise [M] x (K,B,L) are the FULL draws (same on every rank;
    drawn here from a rank-independent generator when absent).  -> {'likelihood': mean over all datapoints}, per [B]."""
    rank, world = rank_world(group)
    B = data[0].shape[0]
    M, L, dev = model.mod, model.params.latent_dim, data[0].device
    if noise is None or mask is None:
        g = torch.Generator(device=dev).manual_seed(20240607 + 1000003 * seed)   # rank-independent
        from .vaes import draw_noise
        if noise is None:
            noise = draw_noise(model.params.dist, (M, K, B, L), dev, generator=g)
        if mask is None:
            mask = torch.bernoulli(torch.full((K, B), 0.5, device=dev), generator=g)
    if not torch.is_tensor(noise):
        noise = torch.stack(list(noise))
    if axis == "batch":
        b0, b1 = shard_bounds(B, rank, world)
        if b1 > b0:
            model.compute_joint_likelihood([t[b0:b1] for t in data], K, batch_size_K, mask=mask[:, b0:b1],
                                           noise=noise[:, :, b0:b1])
            per = model.last_per_datapoint
        else:
            per = torch.empty(0, device=dev)
        per = gather_datapoints(per, B, group)
    elif axis == "samples":
        z, mu, std = model._joint_latents(data, K, batch_size_K, mask, noise)
        k0, k1 = shard_bounds(z.shape[1], rank, world)
        if k1 > k0:
            mx, se = model._is_loglik(z[:, k0:k1].contiguous(), data, range(M), "mixture_half", mu, std, K_total=K,
                                      partial=True)
        else:
            mx = torch.full((B,), -math.inf, device=dev)
            se = torch.zeros(B, device=dev)
        per = finish_lse(*combine_lse(mx, se, group), K)
    else:
        raise ValueError(f"axis must be 'batch' or 'samples', got {axis!r}")
    return {"likelihood": per.mean()}, per


@torch.no_grad()
def conditional_likelihood(model, data, cond_mod, gen_mod, K=1000, batch_size_K=100, axis="batch", noise=None,
                           group=None, seed=0):
    """Multi_VAES.compute_conditional_likelihood (multi_vaes.py:271-318) over all ranks.  noise: full (B,Kused,L)."""
    rank, world = rank_world(group)
    B = data[0].shape[0]
    L, dev = model.params.latent_dim, data[0].device
    Ku = model._used(K, batch_size_K)
    if noise is None:
        from .vaes import draw_noise
        noise = draw_noise(model.vaes[cond_mod].dist, (B, Ku, L), dev,
                           generator=torch.Generator(device=dev).manual_seed(20240608 + 7 * cond_mod + gen_mod
                                                                                    + 1000003 * seed))
    if axis == "batch":
        b0, b1 = shard_bounds(B, rank, world)
        if b1 > b0:
            _, per = model.compute_conditional_likelihood([t[b0:b1] for t in data], cond_mod, gen_mod, K, batch_size_K,
                                                          noise=noise[b0:b1])
            per = per.to(dev)
        else:
            per = torch.empty(0, device=dev)
        per = gather_datapoints(per, B, group)
    elif axis == "samples":
        z = model._cond_latents(cond_mod, data[cond_mod], Ku, noise)[0]
        k0, k1 = shard_bounds(Ku, rank, world)
        if k1 > k0:
            mx, se = model._is_loglik(z[:, k0:k1].contiguous(), data, [gen_mod], None, K_total=K, with_prior=False,
                                      partial=True)
        else:
            mx = torch.full((B,), -math.inf, device=dev)
            se = torch.zeros(B, device=dev)
        per = finish_lse(*combine_lse(mx, se, group), K)
    else:
        raise ValueError(f"axis must be 'batch' or 'samples', got {axis!r}")
    return {f"cond_likelihood_{cond_mod}_{gen_mod}": per.mean()}, per
