"""torchrun --nproc-per-node N tools/nccl_equality.py: the N-rank batch-sharded step (CUDA path + NCCL all-reduce) against the
1-rank step on the concatenated batch: loss and the flat gradient equal to 1e-6, parameters after the update to 1e-5; then
graph replays (all-reduce between the two graphs) leave identical parameters on every rank.  Exit code 0 = equal."""
import faulthandler, os, sys
faulthandler.dump_traceback_later(90, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
from oracle import bivae_oracle as O, synth
from tests.helpers import build_mmvae
from mmvae_b200.train import Trainer

spec = O.spec_mnist_svhn(20, "laplace")
sd = synth.make_state_dict(spec, 1)
Bg, K = 32 * world, 6
x = [t.to(dev) for t in synth.make_inputs(spec, Bg, seed=2)]
rs = np.random.RandomState(3)
noise = [synth.make_noise("laplace", (K, Bg, 20), rs).to(dev) for _ in range(2)]
b0, b1 = rank * 32, (rank + 1) * 32
ok = True
# N-rank step on the shard
tr = Trainer(build_mmvae(spec, sd, device=dev), "m_dreg_looser", K=K, distributed=True)
loss = tr._fwd_bwd([t[b0:b1] for t in x], noise=[n[:, b0:b1].contiguous() for n in noise])
tr._reduce()
g_n = tr.fp.grad.clone()
tot = loss.clone()
dist.all_reduce(tot)
tr._update()
p_n = tr.fp.flat.clone()
# 1-rank step on the whole batch (every rank computes it redundantly)
tr1 = Trainer(build_mmvae(spec, sd, device=dev), "m_dreg_looser", K=K, distributed=False)
loss1 = tr1._fwd_bwd(x, noise=noise)
g_1 = tr1.fp.grad.clone()
tr1._update()
rel = lambda a, b: ((a.double() - b.double()).norm() / b.double().norm()).item()
e_l, e_g, e_p = abs(tot.item() - loss1.item()) / abs(loss1.item()), rel(g_n, g_1), rel(p_n, tr1.fp.flat)
ok &= e_l < 1e-6 and e_g < 1e-6 and e_p < 1e-5      # Adam's first step is g/|g|: entries with |g| ~ 1e-7 of the norm may flip
# graph replay: same parameters on every rank afterwards, and a finite loss
tr.capture([t[b0:b1] for t in x], warmup=2)
for _ in range(3):
    l = tr.step_graph([t[b0:b1] for t in x])
torch.cuda.synchronize()
ref = tr.fp.flat.clone()
dist.broadcast(ref, 0)
same = torch.equal(ref, tr.fp.flat)
ok &= same and bool(torch.isfinite(l))
flag = torch.tensor([int(ok)], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"world {world}: loss rel {e_l:.2e}, flat gradient rel {e_g:.2e}, parameters after update rel {e_p:.2e}; "
          f"graph replay: parameters identical across ranks {same} -> {'OK' if flag.item() else 'MISMATCH'}")
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
