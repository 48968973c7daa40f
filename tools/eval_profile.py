"""Per-C-ABI-call time of one K=1000 evaluator pass (joint + both conditionals, 64 datapoints) of the C2 MMVAE — where the
eval leg of bench.py spends its time.    python tools/eval_profile.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                  # noqa: E402
from mmvae_b200 import ops                    # noqa: E402

dev = torch.device("cuda:0")
model = bench.make_model(dev)
model.branch_streams = False
data = bench.synth_batch(64, seed=999, device=dev)


def once():
    for c, g in ((0, 1), (1, 0)):
        model.compute_conditional_likelihood(data, c, g, K=1000, batch_size_K=100)
    return model.compute_joint_likelihood(data, K=1000, batch_size_K=100)


once()
torch.cuda.synchronize()
t0 = time.perf_counter()
once()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) * 1e3
ops.PROFILE = []
once()
torch.cuda.synchronize()
rec, ops.PROFILE = ops.PROFILE, None
agg = {}
for name, work, e0, e1 in rec:
    a = agg.setdefault(name, [0.0, 0])
    a[0] += e0.elapsed_time(e1)
    a[1] += 1
tot = sum(a[0] for a in agg.values())
print(f"wall {wall:.2f} ms per pass (serial streams); sum of kernel times {tot:.2f} ms over {len(rec)} calls")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:34s} {a[0]:8.3f} ms {a[1]:4d} calls {a[0] / tot * 100:5.1f} %")
