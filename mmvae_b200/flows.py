"""MADE / MAF / IAF normalizing flows with the reference's module layout and state_dict keys
(`flow.net.{b}.net.{0,2,4,6}.{weight,bias,mask}`), so reference checkpoints load.

Reference: my_pythae/models/normalizing_flows/layers.py:14-25 (MaskedLinear), made/made_model.py:21-111 (MADE, the
sequential-degree masks of :68-95), maf/maf_model.py:49-110 (MAF), iaf/iaf_model.py:48-107 (IAF).

The masked contractions run through libmmvae_b200's GEMM (tcgen05 where the shape qualifies) on the pre-masked
weight, which is formed ONCE per flow call and reused by all L steps of the sequential direction (the reference
re-materialises mask*W in every MaskedLinear call: 4 x L x n_blocks times per inverse).  The affine couplings and the
log|det J| accumulation are the fused kernels of csrc/flow.cu.
"""
import torch
import torch.nn as nn

from . import ops
from .networks import ModelOutput


def made_masks(input_dim, hidden_sizes, output_dim):
    """made_model.py:68-95 with degrees_ordering='sequential'."""
    m = {-1: torch.arange(input_dim)}
    for i, h in enumerate(hidden_sizes):
        m[i] = torch.arange(h) % (input_dim - 1)
    masks = [(m[i].unsqueeze(-1) >= m[i - 1].unsqueeze(0)).float() for i in range(len(hidden_sizes))]
    last = (m[len(hidden_sizes) - 1].unsqueeze(0) < m[-1].unsqueeze(-1)).float()
    masks.append(last.repeat(output_dim // input_dim, 1))
    return masks


class MaskedLinear(nn.Linear):
    """layers.py:14-25: a Linear whose weight is multiplied by a fixed 0/1 mask (registered buffer `mask`).
    Parameter container only: the contraction is issued by MADE through ops.LinearAct."""

    def __init__(self, in_features, out_features, mask):
        super().__init__(in_features, out_features)
        self.register_buffer("mask", mask)

    def forward(self, x):
        return ops.LinearAct.apply(x, ops.MaskWeight.apply(self.weight, self.mask), self.bias, None)


class MADE(nn.Module):
    """made_model.py:21-111: [L -> H relu]*n_hidden -> 2L, autoregressive masks; returns (mu, log_var)."""

    def __init__(self, input_dim, hidden_sizes, output_dim=None):
        super().__init__()
        self.input_dim = input_dim
        self.output_dim = 2 * input_dim if output_dim is None else output_dim
        self.hidden_sizes = list(hidden_sizes)
        sizes = [input_dim] + self.hidden_sizes + [self.output_dim]
        masks = made_masks(input_dim, self.hidden_sizes, self.output_dim)
        net = []
        for i, (a, b) in enumerate(zip(sizes[:-1], sizes[1:])):
            net.append(MaskedLinear(a, b, masks[i]))
            if i < len(sizes) - 2:
                net.append(nn.ReLU())
        self.net = nn.Sequential(*net)

    def layers(self):
        return [m for m in self.net if isinstance(m, MaskedLinear)]

    def masked_weights(self):
        """mask*W of every layer, formed once per flow call."""
        return [ops.MaskWeight.apply(l.weight, l.mask) for l in self.layers()]

    def raw(self, x, mw=None):
        """-> h [R, 2L] = (mu | log_var)."""
        lay = self.layers()
        if mw is None and x.is_cuda and ops.made_fused_ok([(l.weight, l.mask, l.bias) for l in lay], x.shape[0]):
            return ops.MadeFused.apply(x, *[t for l in lay for t in (l.weight, l.mask, l.bias)])
        mw = self.masked_weights() if mw is None else mw
        h = x
        for i, (l, w) in enumerate(zip(lay, mw)):
            h = ops.LinearAct.apply(h, w, l.bias, "relu" if i < len(lay) - 1 else None)
        return h

    def forward(self, x):
        h = self.raw(x.reshape(x.shape[0], -1))
        return ModelOutput(mu=h[:, :self.input_dim], log_var=h[:, self.input_dim:])


FUSED_SEQUENTIAL = True      # tests switch it off to compare with the L-pass GEMM formulation


class _Flow(nn.Module):
    def __init__(self, input_dim, n_made_blocks=2, n_hidden_in_made=3, hidden_size=128):
        super().__init__()
        self.input_dim = input_dim
        self.net = nn.ModuleList([MADE(input_dim, [hidden_size] * n_hidden_in_made) for _ in range(n_made_blocks)])

    @staticmethod
    def _zeros(x):
        return torch.zeros(x.shape[0], device=x.device, dtype=torch.float32)

    def _parallel(self, x, mode, blocks, flip_first):
        """One MADE pass per block.  mode 0: u=(x-mu)e^{-lv} then flip (MAF.forward); mode 1: flip then u=x e^{lv}+mu."""
        ladj = self._zeros(x)
        for made in blocks:
            if flip_first:
                x = ops.FlipCols.apply(x)
            x, ladj = ops.FlowAffine.apply(x, made.raw(x), None, ladj, mode, -1, not flip_first)
        return x, ladj

    def _sequential(self, y, mode, blocks, flip_first):
        """L MADE passes per block (maf_model.py:91-104 / iaf_model.py:60-78): column i of the output is produced from
        the MADE of the columns already produced."""
        ladj = self._zeros(y)
        L = self.input_dim
        if FUSED_SEQUENTIAL and not (torch.is_grad_enabled() and (y.requires_grad or any(p.requires_grad for p in self.parameters()))) \
                and all(ops.made_seq_plan(made).ok for made in blocks):
            # inference: one kernel per block, every hidden unit evaluated once (csrc/made_seq.cu)
            for made in blocks:
                y, ladj = ops.made_seq(y, ladj, made, made.masked_weights(), mode, flip_first, not flip_first)
            return y, ladj
        for made in blocks:
            if flip_first:
                y = ops.FlipCols.apply(y)
            mw = made.masked_weights()
            x = torch.zeros_like(y)
            for i in range(L):
                x, ladj = ops.FlowAffine.apply(y, made.raw(x, mw), x, ladj, mode, i, False)
            y = x if flip_first else ops.FlipCols.apply(x)
        return y, ladj


class MAF(_Flow):
    """maf_model.py:23-110.  forward: data -> prior (parallel); inverse: prior -> data (sequential)."""

    def forward(self, x, **kwargs):
        out, ladj = self._parallel(x.reshape(x.shape[0], -1), 0, list(self.net), False)
        return ModelOutput(out=out, log_abs_det_jac=ladj)

    def inverse(self, y, **kwargs):
        out, ladj = self._sequential(y.reshape(y.shape[0], -1), 1, list(self.net)[::-1], True)
        return ModelOutput(out=out, log_abs_det_jac=ladj)


class IAF(_Flow):
    """iaf_model.py:22-107.  forward: the sequential direction; inverse: one pass per block."""

    def forward(self, x, **kwargs):
        out, ladj = self._sequential(x.reshape(x.shape[0], -1), 0, list(self.net), False)
        return ModelOutput(out=out, log_abs_det_jac=ladj)

    def inverse(self, y, **kwargs):
        out, ladj = self._parallel(y.reshape(y.shape[0], -1), 1, list(self.net)[::-1], True)
        return ModelOutput(out=out, log_abs_det_jac=ladj)
