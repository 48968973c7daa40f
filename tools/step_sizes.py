"""C2 training step at several per-GPU batch sizes (strong-scaling shards): eager + graph replay time.  python tools/step_sizes.py 64 32 16"""
import faulthandler, os, sys
faulthandler.dump_traceback_later(100, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mmvae_b200.train import Trainer
dev = torch.device("cuda", 0)
for B in [int(a) for a in sys.argv[1:]] or [128, 64, 32, 16]:
    model = bench.make_model(dev)
    tr = Trainer(model, "m_dreg_looser", K=30, distributed=False)
    xs = [bench.synth_batch(B, seed=5 + i, device=dev) for i in range(4)]
    tr.step(xs[0]); torch.cuda.synchronize()
    print(f"B={B}: eager step ok", flush=True)
    tr.capture(xs[0], warmup=2); torch.cuda.synchronize()
    for i in range(5):
        tr.step_graph(xs[i % 4])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        tr.step_graph(xs[i % 4])
    e1.record(); torch.cuda.synchronize()
    print(f"B={B}: graph replay {e0.elapsed_time(e1) / 20:.3f} ms/step  ({B / (e0.elapsed_time(e1) / 20) * 1e3:.0f} samples/s)", flush=True)
