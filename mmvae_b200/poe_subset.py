"""Importance-sampled conditional likelihoods ln p(x_gen | x_subset) under the product-of-experts posterior of a two-
modality subset — SURVEY.md §8(f) rank 4.  Reference: models/modalities/trimodal.py:158-230
(`compute_cond_ll_from_poe_subset`, `compute_all_cond_ll_from_poe_subsets`), called by MVAE msf.py:101-108.

The reference loops over datapoints and chunks of batch_size_K latents; here all datapoints that share an importance
distribution are processed together: one encoder pass, one rsample launch for (B_s, K_used) latents, one decoder pass
over B_s*K_used rows (chunked by the model's `eval_max_rows`), and the logsumexp kernels of the other evaluators.
"""
import numpy as np
import torch

from . import ops


def _latents_and_logq(model, a, o, data, idx, Ku, noise):
    """z (B_s,Ku,L) ~ q(z|x_a) and ln q(z|x_o) - ln p(z) (B_s,Ku) for the datapoints idx."""
    from .mvae import MVAE
    xa, xo = data[a][idx].contiguous(), data[o][idx].contiguous()
    Bs = xa.shape[0]
    if isinstance(model, MVAE):                        # infer_latent_from_mod, mvae.py:64-70: PoE of q_a and the prior
        mu_u, lv_u = model.vaes[a].encode_raw(xa)
        mu, lnV = model._poe_lnv([mu_u], [lv_u])
        L = mu.shape[1]
        if noise is None:
            noise = torch.randn(Bs, Ku, L, device=mu.device)
        z, _ = ops.RSample.apply(mu.unsqueeze(0), lnV.unsqueeze(0), noise.reshape(1, Bs, Ku, L), "normal", "exp_half", True)
        z = z.reshape(Bs, Ku, L)
        sub_data = [None] * model.mod
        sub_data[o] = xo
        lnq = model.compute_poe_posterior([o], z.reshape(Bs * Ku, L), sub_data)
    else:                                              # multi_vaes.py:71-79: vaes[a](x).z (through the flow)
        z = model._cond_latents(a, xa, Ku, noise)[0]
        L = z.shape[2]
        rows = z.reshape(Bs * Ku, L).contiguous()
        vae = model.vaes[o]
        z0, ladj = vae.flow_forward(rows)
        mu_o, lv_o = vae.encode_raw(xo)
        lnq = ops.GaussRows.apply(z0.contiguous(), mu_o, lv_o, "log_var", Ku)
        if ladj is not None:
            lnq = lnq + ladj
        lnq = lnq + (0.5 * (rows * rows + float(np.log(2 * np.pi)))).sum(dim=1)
    return z.contiguous(), lnq.reshape(Bs, Ku)


def _lse(a):
    """logsumexp over the last dim of (B, n) through the evaluator kernels."""
    mx, se = ops.lse_partial(a.contiguous())
    return ops.lse_finish(mx, se, 1)


@torch.no_grad()
def compute_cond_ll_from_poe_subset(model, data, subset, gen_mod, K=1000, batch_size_K=200, i_s=None, noise=None):
    """trimodal.py:158-216.  For each datapoint one member of `subset` (i_s[b] in {0,1}; np.random.randint when None) is
    the importance distribution, the other enters the weight q(z|x_other)/p(z).  As in the reference, `lnpxs` is re-created
    inside the chunk loop (:188): the numerator is the logsumexp over the LAST chunk of batch_size_K latents, the
    normaliser the logsumexp of the weights over all K // batch_size_K chunks.  noise: (B, K_used, L).
    -> 0-d tensor (mean over datapoints); per-datapoint values in model.last_per_datapoint."""
    nb = K // batch_size_K
    if nb < 1:
        raise ValueError(f"K={K} is smaller than batch_size_K={batch_size_K}: no latent would be drawn")
    Ku = nb * batch_size_K
    B = len(data[0])
    dev = data[0].device
    if i_s is None:
        i_s = np.random.randint(0, 2, size=B)
    i_s = torch.as_tensor(np.asarray(i_s), device=dev)
    ll = torch.empty(B, device=dev, dtype=torch.float32)
    step = max(1, model.eval_max_rows // Ku)
    for v in (0, 1):
        idx_all = torch.nonzero(i_s == v).reshape(-1)
        a, o = subset[v], subset[1 - v]
        for c0 in range(0, idx_all.numel(), step):
            idx = idx_all[c0:c0 + step]
            z, lnq = _latents_and_logq(model, a, o, data, idx, Ku, None if noise is None else noise[idx].contiguous())
            Bs, _, L = z.shape
            lpx = model._lpx_rows(gen_mod, z.reshape(Bs * Ku, L), data[gen_mod][idx].contiguous(), Ku, False).reshape(Bs, Ku)
            num = _lse((lpx + lnq)[:, Ku - batch_size_K:])
            ll[idx] = num - _lse(lnq)
    model.last_per_datapoint = ll
    return torch.sum(ll) / len(ll)


@torch.no_grad()
def compute_all_cond_ll_from_poe_subsets(model, data, K=1000, batch_size_K=100):
    """trimodal.py:221-229 (batch_size_K is accepted and, as in the reference, not forwarded: chunks of 200)."""
    subsets = [[1, 2], [0, 2], [0, 1]]
    r_dict = {}
    for s, gen_mod in zip(subsets, range(3)):
        r_dict[f"cond_poe_ll_{gen_mod}"] = compute_cond_ll_from_poe_subset(model, data, s, gen_mod, K)
    return r_dict
