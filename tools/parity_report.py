"""Parity report (GPU box): for every golden objective case, per gradient tensor, the distance between
   ours (CUDA path; tcgen05 3xTF32 GEMM or exact-fp32 SIMT GEMM), the fp32 REFERENCE digests (tests/golden) and the
   fp64 oracle.  Prints the worst tensor of each (case, objective).  Used to set the tolerances in tests/test_gpu_model.py
   and quoted in DESIGN.md §parity.      python tools/parity_report.py [--all]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from oracle import bivae_oracle as O
from tests.helpers import build_mmvae, leaf_sd, load_golden
from tests.test_gpu_model import OBJ_CASES, _objective
from mmvae_b200 import ops

show_all = "--all" in sys.argv


def sampled_err(f, ref_vals, idx):
    f = f.detach().reshape(-1).double().cpu()
    r = ref_vals.double()
    return ((f[idx] - r).norm() / r.norm().clamp_min(1e-30)).item()


for name in OBJ_CASES:
    g, spec, sd = load_golden(name)
    x = [t.cuda() for t in g["x"]]
    noise = [n.cuda() for n in g["noise"]]
    for oname, res in g["results"].items():
        lsd = leaf_sd(sd, dtype=torch.float64)
        l64 = O.OBJECTIVES[oname](spec, lsd, [t.double() for t in g["x"]], g["K"], [n.double() for n in g["noise"]])
        l64.backward()
        g64 = {k: v.grad for k, v in lsd.items() if v.grad is not None}
        for tc in (True, False):
            ops.ALLOW_TC = tc
            model = build_mmvae(spec, sd)
            loss = _objective(model, oname, x, g["K"], noise)
            loss.backward()
            grads = {k: p.grad for k, p in model.named_parameters()}
            worst = {"o64": (0, ""), "r64": (0, ""), "or": (0, "")}
            tot_o, tot_r = 0.0, 0.0
            for k, d in res["grads"].items():
                if d["norm"] == 0.0:
                    continue
                idx = d["idx"]
                e_o64 = sampled_err(grads[k], g64[k].reshape(-1)[idx], idx)
                e_r64 = ((d["vals"].double() - g64[k].reshape(-1)[idx]).norm() / g64[k].reshape(-1)[idx].norm()).item()
                e_or = sampled_err(grads[k], d["vals"], idx)
                if show_all:
                    print(f"    {k:45s} ours-f64 {e_o64:.2e}  ref32-f64 {e_r64:.2e}  ours-ref32 {e_or:.2e}")
                for key, v in (("o64", e_o64), ("r64", e_r64), ("or", e_or)):
                    if v > worst[key][0]:
                        worst[key] = (v, k)
            keys = sorted(g64)
            ours = torch.cat([grads[k].detach().double().cpu().reshape(-1) for k in keys])
            theirs = torch.cat([g64[k].reshape(-1) for k in keys])
            glob = ((ours - theirs).norm() / theirs.norm()).item()
            print(f"{name:16s} {oname:12s} {'tc3 ' if tc else 'simt'} loss rel(f64) {abs(loss.item() - l64.item()) / abs(l64.item()):.1e} "
                  f"rel(ref32) {abs(loss.item() - res['loss'].item()) / abs(res['loss'].item()):.1e} | global grad vs f64 {glob:.1e} | "
                  f"worst tensor: ours-f64 {worst['o64'][0]:.1e} ({worst['o64'][1]}), ref32-f64 {worst['r64'][0]:.1e} "
                  f"({worst['r64'][1]}), ours-ref32 {worst['or'][0]:.1e}", flush=True)
ops.ALLOW_TC = True
