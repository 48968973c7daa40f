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
