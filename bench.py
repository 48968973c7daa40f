"""bench.py — MNIST-SVHN MMVAE K=30 DReG training samples/s (+ K=1000 NLL eval samples/s) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W]                 # own arm (CUDA path through the C ABI)
    python bench.py --impl reference [--gpus N] --steps K --warmup W    # reference arm: the CPU port on host cores

Workload = BASELINE.json configs[1] (C2): M=2, x0 (B,1,28,28), x1 (B,3,32,32), L=20, Laplace posterior/prior,
Normal(sigma=1) likelihood, lik_scaling=(3072/784,1), objective m_dreg_looser, K=30, B=128 per GPU (weak scaling),
Adam(amsgrad, 1e-3).  A "step" = forward + backward + gradient all-reduce (N>1) + optimizer step on one synthetic
batch.  One JSON line is printed by rank 0 (contract in the task statement).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

_OUT = sys.stdout
ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "MNIST-SVHN MMVAE K=30 DReG train samples/s"
B_PER_GPU, K_TRAIN, LATENT = 128, 30, 20
FLOP_PER_SAMPLE = 1.845e9          # necessary fwd+bwd work per training sample at C2
EVAL_FLOP_PER_DATAPOINT = 20.2e9   # joint + both conditionals at K=1000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-eval", action="store_true", help="skip the K=1000 evaluator leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eval-datapoints", type=int, default=64)
    ap.add_argument("--detail", action="store_true", help="print every C-ABI call of the profiled step to stderr")
    return ap.parse_args()


TC_FAMILY = ("mmvae_gemm", "mmvae_convt_tc_fwd", "mmvae_convt_tc_dgrad", "mmvae_convt_tc_wgrad")


def load_traffic():
    """mean ncu dram bytes (read + write) per launch of a kernel, from the committed `ncu --set full` summaries."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p))
    except (OSError, ValueError):
        return {}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "which": "measured"}
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "which": "fallback"}


# ------------------------------------------------------------------------------------------------ reference arm
def cpu_port_step_rate(steps, warmup, B=B_PER_GPU, K=K_TRAIN):
    """The CPU port (oracle/bivae_oracle.py, pinned to the reference by tests/golden) timed on the host cores:
    m_dreg_looser forward + backward + Adam(amsgrad) at the C2 shapes.  Returns (samples/s, cores, ms/step)."""
    from oracle import bivae_oracle as O
    from oracle import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = O.spec_mnist_svhn(LATENT, "laplace")
    sd = synth.make_state_dict(spec, 1)
    leaves = {k: v.clone().requires_grad_(k.endswith(("weight", "bias"))) for k, v in sd.items()}
    opt = torch.optim.Adam([v for v in leaves.values() if v.requires_grad], lr=1e-3, amsgrad=True)
    x = synth.make_inputs(spec, B, seed=2)
    rs = np.random.RandomState(3)
    times = []
    for i in range(warmup + steps):
        noise = [synth.make_noise("laplace", (K, B, LATENT), rs) for _ in range(2)]
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        loss = O.m_dreg_looser(spec, leaves, x, K, noise)
        (-loss).backward()
        opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    return B / dt, cores, dt * 1e3


def port_eager_gpu_rate(dev, steps=5, warmup=3, B=B_PER_GPU, K=K_TRAIN):
    """The same port (torch ops, fp32, the reference's op sequence incl. its K-fold encoder-free restatement) run
    EAGERLY on the B200 — the reference's real deployment mode is `model.to('cuda')` (main.py:64-71), so this, not the
    CPU figure, is the number the kernels have to beat.  Reported inside cpu_baseline as `eager_b200`."""
    from oracle import bivae_oracle as O
    from oracle import synth
    spec = O.spec_mnist_svhn(LATENT, "laplace")
    sd = {k: v.to(dev) for k, v in synth.make_state_dict(spec, 1).items()}
    leaves = {k: v.clone().requires_grad_(k.endswith(("weight", "bias"))) for k, v in sd.items()}
    opt = torch.optim.Adam([v for v in leaves.values() if v.requires_grad], lr=1e-3, amsgrad=True)
    x = [t.to(dev) for t in synth.make_inputs(spec, B, seed=2)]
    rs = np.random.RandomState(3)
    noises = [[synth.make_noise("laplace", (K, B, LATENT), rs).to(dev) for _ in range(2)] for _ in range(2)]

    def step(i):
        opt.zero_grad(set_to_none=True)
        loss = O.m_dreg_looser(spec, leaves, x, K, noises[i % 2])
        (-loss).backward()
        opt.step()
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return B / (ms * 1e-3), ms


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, cores, ms = cpu_port_step_rate(max(1, args.steps), max(0, args.warmup))
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2: MNIST-SVHN MMVAE m_dreg_looser K=30 B=128 L=20 laplace, Adam(amsgrad)",
                   "note": "the reference is pure Python/PyTorch; its CPU path is timed through the CPU port "
                           "(oracle/bivae_oracle.py, pinned to the unmodified reference by tests/golden) on all host "
                           "cores; it does not scale with --gpus"},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} full C2 steps (B=128,K=30) fwd+bwd+Adam after {args.warmup} warm-up"},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()


# ------------------------------------------------------------------------------------------------------ helpers
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_model(device):
    import mmvae_b200
    params = argparse.Namespace(latent_dim=LATENT, dist="laplace", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                device=str(device), data_path="", epochs=1)
    torch.manual_seed(1)                          # main.py:49-51 seeds with args.seed = 1
    return mmvae_b200.VAE_mnist_svhn(params).to(device)


def synth_batch(B, seed, device=None, pin=False):
    """U(0,1) images of the reference loader shapes."""
    g = torch.Generator().manual_seed(seed)
    xs = [torch.rand(B, 1, 28, 28, generator=g), torch.rand(B, 3, 32, 32, generator=g)]
    if pin:
        xs = [t.pin_memory() for t in xs]
    if device is not None:
        xs = [t.to(device) for t in xs]
    return xs


def timed(fn, steps, dist_on):
    """K calls of fn bracketed by barrier + synchronize; CUDA events on the launching stream; max over ranks (ms)."""
    import torch.distributed as dist
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if dist_on:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


def profile_step(trainer, x, peak, detail=False):
    """One instrumented eager step: CUDA events around every C-ABI call -> per-kernel time shares and the roofline
    figures of the dominant kernel (algorithmic work / measured time)."""
    from mmvae_b200 import ops
    branch = getattr(trainer.model, "branch_streams", False)
    trainer.model.branch_streams = False           # serial launches: the per-call CUDA events must not overlap
    wstream, trainer._wgrad_stream = trainer._wgrad_stream, None
    for _ in range(2):
        trainer.step(x)
    torch.cuda.synchronize()
    ops.PROFILE = []
    ops.GEMM_LOG = [] if detail else None
    n0 = ops.N_CALLS
    trainer.step(x)
    torch.cuda.synchronize()
    rec, ops.PROFILE = ops.PROFILE, None
    trainer.model.branch_streams = branch
    trainer._wgrad_stream = wstream
    calls = ops.N_CALLS - n0
    agg = {}
    if detail:
        for name, work, e0, e1 in rec:
            sys.stderr.write(f"{name:32s} {e0.elapsed_time(e1) * 1e3:9.1f} us  {work}\n")
        for tag, e0, e1 in ops.GEMM_LOG or []:
            sys.stderr.write(f"gemm {tag}  {e0.elapsed_time(e1) * 1e3:9.1f} us\n")
        ops.GEMM_LOG = None
    for name, work, e0, e1 in rec:
        a = agg.setdefault(name, {"ms": 0.0, "n": 0, "flop": 0.0, "byte": 0.0})
        a["ms"] += e0.elapsed_time(e1)
        a["n"] += 1
        if work is not None:
            a[work[0]] += work[1]
    total = sum(a["ms"] for a in agg.values())
    shares = {k: {"ms": round(a["ms"], 4), "launches": a["n"], "share": round(a["ms"] / total, 4)}
              for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"])}
    # the dominant KERNEL: gemm_tc_kernel (tcgen05 3xTF32) is launched through four C-ABI entry points
    fam = {"ms": 0.0, "n": 0, "flop": 0.0}
    for k in TC_FAMILY:
        if k in agg:
            for f in fam:
                fam[f] += agg[k][f]
    rest = {k: a for k, a in agg.items() if k not in TC_FAMILY}
    top_rest = max(rest, key=lambda k: rest[k]["ms"]) if rest else None
    traffic = load_traffic()
    if fam["ms"] >= (rest[top_rest]["ms"] if top_rest else 0.0):
        ach = fam["flop"] / (fam["ms"] * 1e-3) / 1e12
        roof = {"kernel": "gemm_tc_kernel (tcgen05 3xTF32 GEMM / implicit-GEMM convT; entry points " + ", ".join(
                    k for k in TC_FAMILY if k in agg) + ")",
                "bound": "tensor", "achieved": ach, "peak": peak["tflops"], "unit": "TFLOP/s",
                "frac": ach / peak["tflops"], "traffic": traffic.get("gemm_tc_kernel"),
                "launches": fam["n"], "share_of_step": round(fam["ms"] / total, 4),
                "tensor_pipe_frac": 6.0 * ach / peak["tflops"],
                "note": f"algorithmic 2MNK of the {fam['n']} launches of one step / their summed CUDA-event time; peak = "
                        f"{peak['which']} dense bf16 (sustained). The path computes fp32-accurate 3xTF32: 3 tf32 MMAs per "
                        "algorithmic MMA on a pipe whose tf32 rate is half the bf16 rate, so tensor_pipe_frac = 6 x frac; "
                        "traffic = mean ncu dram bytes per launch (profiles/traffic.json) or null"}
    else:
        a = rest[top_rest]
        ach = a["byte"] / (a["ms"] * 1e-3) / 1e9
        roof = {"kernel": top_rest, "bound": "hbm", "achieved": ach, "peak": peak["hbm_gbs"], "unit": "GB/s",
                "frac": ach / peak["hbm_gbs"], "traffic": traffic.get(top_rest), "launches": a["n"],
                "share_of_step": round(a["ms"] / total, 4),
                "note": f"algorithmic bytes of the {a['n']} {top_rest} launches of one step / their summed CUDA-event "
                        f"time; peak = {peak['which']} copy bandwidth"}
    return roof, shares, calls, total


# --------------------------------------------------------------------------------------------------------- main
def run_b200(args):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (own arm) needs a CUDA device: mmvae_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist_on = world > 1
    if dist_on:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")        # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    from mmvae_b200 import ops
    from mmvae_b200.train import Trainer
    peak = peaks()
    model = make_model(dev)
    trainer = Trainer(model, "m_dreg_looser", K=K_TRAIN, lr=1e-3, amsgrad=True)
    trainer.broadcast_parameters(0)
    B = B_PER_GPU
    batches = [synth_batch(B, seed=100 + rank * 17 + i, device=dev) for i in range(4)]
    host_batches = [synth_batch(B, seed=100 + rank * 17 + i, pin=True) for i in range(4)]

    # launches of one eager step (also the first compile/warm-up of every kernel)
    n0 = ops.N_CALLS
    trainer.step(batches[0])
    torch.cuda.synchronize()
    calls_per_step = ops.N_CALLS - n0

    use_graph = not args.no_graph
    graph_err = None
    if use_graph:
        try:
            trainer.capture(batches[0], warmup=2)
        except Exception as e:                      # same kernels, launched eagerly instead of replayed
            graph_err = f"{type(e).__name__}: {e}"[:200]
            use_graph = False
            torch.cuda.synchronize()
    step_dev = (lambda i: trainer.step_graph(batches[i % 4])) if use_graph else (lambda i: trainer.step(batches[i % 4]))

    def step_e2e(i):
        hb = host_batches[i % 4]
        if use_graph:                                                   # H2D of this step's inputs (pinned) ...
            loss = trainer.step_graph(hb)                               # ... straight into the graph's input buffers
        else:
            loss = trainer.step([t.to(dev, non_blocking=True) for t in hb])
        return loss.item()                                              # D2H of the step's result (main.py:160)

    for i in range(max(3, args.warmup)):
        step_dev(i)
    clocks = ClockSampler(local).start()
    ms = timed(step_dev, args.steps, dist_on)
    ms_per_step = ms / args.steps
    value = world * B / (ms_per_step * 1e-3)
    for i in range(3):
        step_e2e(i)
    ms_e2e = timed(step_e2e, args.steps, dist_on) / args.steps
    clk = clocks.stop()
    e2e_value = world * B / (ms_e2e * 1e-3)
    h2d = sum(t.numel() * 4 for t in host_batches[0])

    line = {
        "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2: MNIST-SVHN MMVAE m_dreg_looser K=30 L=20 laplace posterior, Normal likelihood, "
                               "Adam(amsgrad) step included", "batch_per_gpu": B, "global_batch": B * world, "K": K_TRAIN,
                   "parallelism": f"dp{world} batch-sharded, one flat NCCL all-reduce(SUM) of 1289507 fp32 gradients",
                   "cuda_graph": use_graph, "graph_error": graph_err,
                   "l2": "per-step working set (decoder activations of 7680 latent rows, ~2 GB) exceeds the 126 MB L2; "
                         "4 input batches rotate"},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e},
        "gpu_launches": calls_per_step * args.steps,
        "achieved_tflops_algorithmic": value * FLOP_PER_SAMPLE / 1e12 / world,
    }
    prof = profile_step(trainer, batches[0], peak, args.detail)      # every rank: the step contains the all-reduce
    if rank == 0:
        roof, shares, calls, prof_ms = prof
        line["roofline"] = roof
        line["kernel_shares"] = shares
        line["profiled_step_ms"] = prof_ms
    ev = None
    if not args.no_eval:
        ev = run_eval(model, dev, args.eval_datapoints * world, peak, dist_on)       # every rank: batch-sharded
        ev_jnf = run_eval_jmvae_nf(dev, args.eval_datapoints, dist_on)
    if rank == 0:
        if ev is not None:
            line["eval"] = ev
            line["eval_jmvae_nf"] = ev_jnf
        if world == 1 and not args.no_cpu_baseline:
            v, cores, cms = cpu_port_step_rate(2, 1)
            line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
                                    "sample": "2 full C2 steps (B=128,K=30) fwd+bwd+Adam after 1 warm-up",
                                    "ms_per_step": cms}
            try:
                gv, gms = port_eager_gpu_rate(dev)
                line["cpu_baseline"]["eager_b200"] = {
                    "value": gv, "unit": "samples/s", "ms_per_step": gms,
                    "what": "the same port (torch eager ops, cuDNN/cuBLAS, TF32 off) on this B200: the reference's own "
                            "deployment mode; 5 steps after 3 warm-up"}
            except Exception as e:                  # informational only
                line["cpu_baseline"]["eager_b200"] = {"error": f"{type(e).__name__}: {e}"[:200]}
        _OUT.write(json.dumps(line) + "\n")
        _OUT.flush()
    if dist_on:
        dist.barrier()
        dist.destroy_process_group()


def run_eval(model, dev, n, peak, dist_on=False):
    """K=1000 NLL evaluator (compute_likelihoods.py:106-107): joint + both conditionals per datapoint; n datapoints in
    total, sharded over the ranks by batch (mmvae_b200.parallel)."""
    import torch.distributed as dist
    from mmvae_b200 import parallel as P
    data = synth_batch(n, seed=999, device=dev)

    def once():
        for c, g in ((0, 1), (1, 0)):
            P.conditional_likelihood(model, data, c, g, K=1000, batch_size_K=100, axis="batch")
        return P.joint_likelihood(model, data, K=1000, batch_size_K=100, axis="batch")[0]["likelihood"]
    once()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 2
    for _ in range(reps):
        ll = once()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if dist_on:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    s = ms.item() * 1e-3 / reps
    return {"metric": "K=1000 NLL eval samples/s (joint + both conditionals)", "value": n / s, "unit": "samples/s",
            "datapoints": n, "K": 1000, "batch_size_K": 100, "s_per_pass": s, "joint_ll": float(ll),
            "sharding": "batch", "achieved_tflops_algorithmic": n * EVAL_FLOP_PER_DATAPOINT / s / 1e12}


def run_eval_jmvae_nf(dev, n_per_rank, dist_on=False):
    """The same K=1000 evaluator trio on BASELINE config C3's model (JMVAE-NF, MAF flows, DCCA two-step encoders,
    jmvae_nf_dcca.json): joint ln p(x,y) from the joint encoder and both conditionals, whose latents go through the
    sequential direction of the flow (csrc/made_seq.cu).  Each rank evaluates its own n_per_rank datapoints."""
    import argparse
    import torch.distributed as dist
    import mmvae_b200
    params = argparse.Namespace(latent_dim=20, dist="normal", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                num_hidden_layers=1, beta_kl=1.0, decrease_beta_kl=1.0, fix_jencoder=True, fix_decoders=True,
                                no_recon=True, warmup=100, dcca=True, dim_dcca=9, flow="maf", no_nf=False, device="cuda",
                                data_path="", epochs=1)
    torch.manual_seed(1)
    model = mmvae_b200.VAE_jnf_mnist_svhn_dcca(params).to(dev)
    data = synth_batch(n_per_rank, seed=998, device=dev)

    def once():
        for c, g in ((0, 1), (1, 0)):
            model.compute_conditional_likelihood(data, c, g, K=1000, batch_size_K=100)
        return model.compute_joint_likelihood(data, K=1000, batch_size_K=100)["likelihood"]
    once()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 2
    for _ in range(reps):
        ll = once()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    world = 1
    if dist_on:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        world = dist.get_world_size()
    s = ms.item() * 1e-3 / reps
    return {"metric": "K=1000 NLL eval samples/s (joint + both conditionals), JMVAE-NF MAF model (config C3)",
            "value": n_per_rank * world / s, "unit": "samples/s", "datapoints": n_per_rank * world, "K": 1000,
            "batch_size_K": 100, "s_per_pass": s, "joint_ll": float(ll), "sharding": "batch"}


def _claim_stdout():
    """Libraries (NCCL's version banner, ...) write to fd 1; the contract is ONE JSON line on stdout.  Point fd 1 at
    stderr for the duration of the run and keep the real stdout for the final line."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    a = parse()
    _OUT = _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
