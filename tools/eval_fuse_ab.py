"""A/B of the evaluators with the last decoder layer fused with the pixel log-likelihood (Decoder_VAE_SVHN.fuse_eval_loglik).  python tools/eval_fuse_ab.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = sys.argv[:1]
import torch
import bench as Bn
dev = torch.device("cuda", 0)
model = Bn.make_model(dev)
for fuse in (False, True, False, True):
    model.vaes[1].decoder.fuse_eval_loglik = fuse
    ev = Bn.run_eval(model, dev, 500, Bn.peaks())
    print(f"fuse_eval_loglik={fuse}: {ev['value']:.0f} datapoints/s ({ev['s_per_pass']*1e3:.2f} ms/pass) joint_ll {ev['joint_ll']:.3f}", flush=True)
