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
EVAL_METRIC = "K=1000 NLL eval samples/s (joint + both conditionals)"
B_PER_GPU, K_TRAIN, LATENT = 128, 30, 20
FLOP_PER_SAMPLE = 1.845e9          # necessary fwd+bwd work per training sample at C2
EVAL_FLOP_PER_DATAPOINT = 20.2e9   # joint + both conditionals at K=1000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: B=128 per GPU (default, the driver's contract); strong: the reference config's GLOBAL B=128 split over the GPUs")
    ap.add_argument("--one-graph", action="store_true", help="capture the NCCL all-reduce inside the step graph (opt-in, see train.py)")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-eval", action="store_true", help="skip the K=1000 evaluator leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eval-datapoints", type=int, default=500, help="per rank; C4 = the loader batch of compute_likelihoods.py:82")
    ap.add_argument("--detail", action="store_true", help="print every C-ABI call of the profiled step to stderr")
    return ap.parse_args()


# C-ABI entry point -> the kernel it launches (the roofline is reported per KERNEL; mmvae_gemm is a dispatcher and is split by
# the kernel it picked, see ops.gemm)
KERNEL_OF = {
    "mmvae_convt_ts_fwd": "convt_ts_kernel<FWD>", "mmvae_convt_ts_dgrad": "convt_ts_kernel<DGRAD>",
    "mmvae_convt_ts_wgrad": "convt_ts_wgrad_kernel", "mmvae_gemm[tcgen05]": "gemm_tc_kernel", "mmvae_gemm[simt]": "gemm_simt_kernel",
    "mmvae_gemm[skinny]": "gemm_skinny_kernel", "mmvae_gemm[small]": "gemm_small_kernel",
    "mmvae_convt_tc_fwd": "convt_fwd9_kernel", "mmvae_convt_tc_dgrad": "gemm_tc_kernel(dgrad)", "mmvae_convt_tc_wgrad": "gemm_tc_kernel(wgrad)",
    "mmvae_convt_small_fwd": "convt_l4_fwd_tc_kernel", "mmvae_convt_small_dgrad": "convt_l4_dgrad_tc_kernel",
    "mmvae_convt_small_wgrad": "convt_l4w_ts_kernel", "mmvae_loglik_fwd": "loglik_fwd_kernel", "mmvae_loglik_bwd": "loglik_bwd_kernel",
}


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


def port_eager_gpu_rate(dev, strict_fp32, steps=5, warmup=3, B=B_PER_GPU, K=K_TRAIN):
    """The CPU port's op sequence (torch ops; it already encodes once instead of K times, so it is FASTER than the reference's
    own forward) run eagerly on the B200 through cuDNN / cuBLAS -- the reference's deployment mode is `model.to('cuda')` with
    `cudnn.benchmark = True` (main.py:47,64-71).  strict_fp32 = False: torch's defaults as the reference runs them (cuDNN
    convolutions may use TF32, matmuls fp32); strict_fp32 = True: TF32 off everywhere, the accuracy class this repo's kernels
    are held to.  Returns (samples/s, ms/step)."""
    from oracle import bivae_oracle as O
    from oracle import synth
    saved = (torch.backends.cudnn.benchmark, torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = not strict_fp32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
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
    finally:
        torch.backends.cudnn.benchmark, torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved
    return B / (ms * 1e-3), ms


def cpu_port_eval_rate(n=4, K=1000, bs=100):
    """The CPU port's three K-sample evaluators (oracle/bivae_oracle.py: joint + both conditionals, the calls of
    compute_likelihoods.py:106-107) on n datapoints on the host cores.  Returns (datapoints/s, cores, seconds)."""
    from oracle import bivae_oracle as O
    from oracle import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = O.spec_mnist_svhn(LATENT, "laplace")
    sd = synth.make_state_dict(spec, 1)
    data = synth.make_inputs(spec, n, seed=4)
    rs = np.random.RandomState(5)
    nz = lambda *shape: synth.make_noise("laplace", shape, rs)
    mask = torch.from_numpy(rs.binomial(1, 0.5, (K, n))).float()
    with torch.no_grad():
        t0 = time.perf_counter()
        O.conditional_likelihood(spec, sd, data, 0, 1, K, bs, nz(n, K, LATENT))
        O.conditional_likelihood(spec, sd, data, 1, 0, K, bs, nz(n, K, LATENT))
        O.mmvae_joint_likelihood(spec, sd, data, K, bs, mask, [nz(K, n, LATENT), nz(K, n, LATENT)])
        dt = time.perf_counter() - t0
    return n / dt, cores, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, cores, ms = cpu_port_step_rate(max(1, args.steps), max(0, args.warmup))
    ev, _, ev_s = cpu_port_eval_rate(4)
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
        "eval": {"metric": EVAL_METRIC, "value": ev, "unit": "samples/s", "datapoints": 4, "K": 1000, "batch_size_K": 100,
                 "s_per_pass": ev_s, "cores": cores, "kind": "port",
                 "sample": "4 datapoints x (joint + both conditionals) at K=1000 on the host cores"},
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
                                          "--format=csv,noheader,nounits", "-lms", "20"],
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


def roofline_of(agg, total, peak, top_n=6):
    """Roofline objects of the KERNELS of a profiled pass (entry points that launch one kernel), largest summed CUDA-event time
    first; the first one is the line's `roofline`, the list its `rooflines`.  achieved = ALGORITHMIC flops (2MNK of the
    contraction; tensor-bound kernels) or bytes (HBM-bound ones) of its launches / their summed time; peak =
    MEASURED_PEAKS.json (sustained bf16 / copy)."""
    traffic = load_traffic()
    cand = {k: a for k, a in agg.items() if k in KERNEL_OF and (a["flop"] > 0 or a["byte"] > 0)}
    out = []
    for top in sorted(cand, key=lambda k: -cand[k]["ms"])[:top_n]:
        a, kern = cand[top], KERNEL_OF[top]
        tr = traffic.get(kern, {}) if isinstance(traffic.get(kern), dict) else {}
        if a["flop"] > 0:
            ach = a["flop"] / (a["ms"] * 1e-3) / 1e12
            roof = {"kernel": kern, "entry_point": top, "bound": "tensor", "achieved": ach, "peak": peak["tflops"], "unit": "TFLOP/s",
                    "frac": ach / peak["tflops"]}
            how = {"gemm_tc_kernel": "3xTF32: three tf32 MMAs per algorithmic MMA, and tf32 runs at half the bf16 rate, so its "
                                     "MMA-work fraction is 6 x frac",
                   "gemm_simt_kernel": "exact fp32 on the FMA pipes (no tensor cores): the tensor peak is only the common scale",
                   "gemm_small_kernel": "exact fp32 on the FMA pipes, latency-bound sizes: the tensor peak is only the common scale",
                   "gemm_skinny_kernel": "K <= 32: store-bound, the tensor peak is only the common scale"}.get(
                kern, "bf16x3: hi*hi + hi*mid + mid*hi on bf16 pieces (3 MMAs per algorithmic MMA), so its MMA-work fraction is 3 x frac")
            note = ("algorithmic 2MNK of the kernel's launches in one pass / their summed CUDA-event time; peak = "
                    f"{peak['which']} dense bf16 (sustained). " + how)
        else:
            ach = a["byte"] / (a["ms"] * 1e-3) / 1e9
            roof = {"kernel": kern, "entry_point": top, "bound": "hbm", "achieved": ach, "peak": peak["hbm_gbs"], "unit": "GB/s",
                    "frac": ach / peak["hbm_gbs"]}
            note = f"algorithmic bytes of the kernel's launches in one pass / their summed CUDA-event time; peak = {peak['which']} copy bandwidth"
        roof.update({"traffic": tr.get("dram_bytes_per_launch"), "launches": a["n"], "us_per_launch": round(1e3 * a["ms"] / a["n"], 1),
                     "share_of_pass": round(a["ms"] / total, 4), "tensor_pipe_active_ncu_pct": tr.get("tensor_pipe_active_pct"),
                     "note": note + "; traffic / tensor_pipe_active_ncu_pct: ncu --set full of the same kernel (profiles/traffic.json), null if absent"})
        out.append(roof)
    return out


def profile_pass(fn):
    """Runs fn() with CUDA events around every C-ABI call -> ({entry: {ms, n, flop, byte}}, total ms, records)."""
    from mmvae_b200 import ops
    torch.cuda.synchronize()
    ops.PROFILE = []
    try:
        fn()
        torch.cuda.synchronize()
    finally:
        rec, ops.PROFILE = ops.PROFILE, None
    agg = {}
    for name, work, e0, e1 in rec:
        a = agg.setdefault(name, {"ms": 0.0, "n": 0, "flop": 0.0, "byte": 0.0})
        a["ms"] += e0.elapsed_time(e1)
        a["n"] += 1
        if work is not None:
            a[work[0]] += work[1]
    return agg, sum(a["ms"] for a in agg.values()), rec


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
    roof = roofline_of(agg, total, peak)
    return roof, shares, calls, total


# --------------------------------------------------------------------------------------------------------- main
def other_config_steps(dev, reps=10):
    """Training step (forward + backward + AMSGrad, CUDA-graph replay through train.Trainer) of the other BASELINE configs on
    one GPU: extra keys of the line, not the metric.  C1 circles-squares m_iwae K=10 B=64 (mmvae_circles.py defaults), C3
    JMVAE-NF MAF + DCCA m_jmvae_nf B=128 (decoders and joint encoder frozen: its second training phase), C5 trimodal
    m_dreg_looser K=30 B=64."""
    import mmvae_b200
    from mmvae_b200 import ops
    from mmvae_b200.train import Trainer
    out = {}

    def ns(**kw):
        base = dict(latent_dim=LATENT, dist="laplace", recon_losses=["normal", "normal"], llik_scaling=0.0, device=str(dev),
                    data_path="", epochs=1)
        base.update(kw)
        return argparse.Namespace(**base)
    cases = {
        "C1 circles-squares m_iwae K=10 B=64 L=4": (lambda: mmvae_b200.VAE_circles_squares(ns(latent_dim=4, dist="normal")), "m_iwae", 10,
                                                [(64, 1, 32, 32), (64, 1, 32, 32)]),
        "C3 JMVAE-NF MAF+DCCA m_jmvae_nf B=128 (flows phase)": (
            lambda: mmvae_b200.VAE_jnf_mnist_svhn_dcca(ns(dist="normal", num_hidden_layers=1, beta_kl=1.0, decrease_beta_kl=1.0,
                                                          fix_jencoder=True, fix_decoders=True, no_recon=True, warmup=0, dcca=True,
                                                          dim_dcca=9, flow="maf", no_nf=False)), "m_jmvae_nf", 1,
            [(128, 1, 28, 28), (128, 3, 32, 32)]),
        "C5 trimodal MNIST-SVHN-Fashion m_dreg_looser K=30 B=64": (
            lambda: mmvae_b200.VAE_mmvae_msf(ns(latent_dim=30, dist="normal", recon_losses=["normal"] * 3)), "m_dreg_looser", 30,
            [(64, 1, 28, 28), (64, 3, 32, 32), (64, 1, 28, 28)]),
    }
    for name, (build, objective, K, shapes) in cases.items():
        try:
            torch.manual_seed(1)
            model = build().to(dev)
            tr = Trainer(model, objective, K=K, lr=1e-3, distributed=False)
            g = torch.Generator().manual_seed(3)
            x = [torch.rand(*s, generator=g).to(dev) for s in shapes]
            n0 = ops.N_CALLS
            tr.step(x)
            torch.cuda.synchronize()
            calls = ops.N_CALLS - n0
            tr.capture(x, warmup=2)
            ms = timed(lambda i: tr.step_graph(x), reps, False) / reps
            out[name] = {"ms_per_step": ms, "samples_per_s": shapes[0][0] / (ms * 1e-3), "launches_per_step": calls}
            del tr, model
        except Exception as e:                              # informational keys only
            out[name] = {"error": f"{type(e).__name__}: {e}"[:200]}
    return out


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
    B = B_PER_GPU if args.scaling == "weak" else max(1, B_PER_GPU // world)
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
            trainer.capture(batches[0], warmup=2, capture_collective=args.one_graph)
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
        "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2: MNIST-SVHN MMVAE m_dreg_looser K=30 L=20 laplace posterior, Normal likelihood, "
                               "Adam(amsgrad) step included", "batch_per_gpu": B, "global_batch": B * world, "K": K_TRAIN,
                   "parallelism": f"dp{world} batch-sharded, one flat NCCL all-reduce(SUM) of 1289507 fp32 gradients",
                   "cuda_graph": use_graph, "graph_error": graph_err,
                   "collective_in_graph": bool(use_graph and dist_on and args.one_graph),
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
        roofs, shares, calls, prof_ms = prof
        line["roofline"] = roofs[0] if roofs else None
        line["rooflines"] = roofs
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
        if world == 1 and not args.no_eval:
            line["other_configs"] = other_config_steps(dev)
        if world == 1 and not args.no_cpu_baseline:
            v, cores, cms = cpu_port_step_rate(2, 1)
            line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
                                    "sample": "2 full C2 steps (B=128,K=30) fwd+bwd+Adam after 1 warm-up",
                                    "ms_per_step": cms}
            try:
                gv, gms = port_eager_gpu_rate(dev, strict_fp32=False)
                sv, sms = port_eager_gpu_rate(dev, strict_fp32=True)
                line["cpu_baseline"]["eager_b200"] = {
                    "reference_settings": {"value": gv, "unit": "samples/s", "ms_per_step": gms,
                                           "what": "cudnn.benchmark=True, torch's default TF32 policy (cuDNN may use TF32, "
                                                   "matmuls fp32): how main.py:47,64-71 runs on a GPU"},
                    "strict_fp32": {"value": sv, "unit": "samples/s", "ms_per_step": sms,
                                    "what": "cudnn.benchmark=True, TF32 disabled for cuDNN and cuBLAS: the accuracy class of "
                                            "this repo's kernels"},
                    "what": "the CPU port's op sequence (it encodes once, i.e. it is faster than the reference's own "
                            "K-tiled forward) run eagerly with torch/cuDNN/cuBLAS on this B200; 5 steps after 3 warm-up"}
            except Exception as e:                  # informational only
                line["cpu_baseline"]["eager_b200"] = {"error": f"{type(e).__name__}: {e}"[:200]}
            if ev is not None:
                ev_v, ev_c, ev_s = cpu_port_eval_rate(4)
                line["eval"]["cpu_baseline"] = {"value": ev_v, "unit": "samples/s", "cores": ev_c, "kind": "port",
                                                "sample": "4 datapoints x (joint + both conditionals) at K=1000", "seconds": ev_s}
        _OUT.write(json.dumps(line) + "\n")
        _OUT.flush()
    if dist_on:
        dist.barrier()
        dist.destroy_process_group()


def run_eval(model, dev, n, peak, dist_on=False):
    """K=1000 NLL evaluator (compute_likelihoods.py:82,106-107: batches of 500 datapoints, K=1000, batch_size_K=100): joint +
    both conditionals per datapoint; n datapoints in total.  Timed (a) batch-sharded over the ranks with the data resident in
    HBM (`value`), (b) the same through host buffers (`e2e`: H2D of the batch inside the timed region, per-datapoint results
    read back), (c) sharded over the K importance samples with the (max, sum-exp) combine (`samples_axis`; equal to (a) on one
    GPU).  Every rank runs it; rank 0 reports."""
    import torch.distributed as dist
    from mmvae_b200 import parallel as P
    data = synth_batch(n, seed=999, device=dev)
    host = synth_batch(n, seed=999, pin=True)

    def once(axis="batch", src=None):
        d = data if src is None else [t.to(dev, non_blocking=True) for t in src]
        for c, g in ((0, 1), (1, 0)):
            _, per = P.conditional_likelihood(model, d, c, g, K=1000, batch_size_K=100, axis=axis)
        out, perj = P.joint_likelihood(model, d, K=1000, batch_size_K=100, axis=axis)
        if src is not None:
            perj.cpu()                                      # the per-datapoint log-likelihoods go back to the host
        return out["likelihood"]

    def clock(fn, reps=2):
        fn()
        if dist_on:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ll = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if dist_on:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() * 1e-3 / reps, float(ll)

    s, ll = clock(once)
    s_e2e, _ = clock(lambda: once(src=host))
    s_k, ll_k = clock(lambda: once(axis="samples")) if dist_on else (s, ll)
    ev = {"metric": EVAL_METRIC, "value": n / s, "unit": "samples/s", "datapoints": n, "K": 1000, "batch_size_K": 100,
          "s_per_pass": s, "joint_ll": ll, "sharding": "batch",
          "config": {"workload": "C4: K=1000 importance-sampled joint + conditional log-likelihoods, MNIST-SVHN MMVAE, "
                                 f"{n} datapoints (loader batch 500 per GPU), batch_size_K=100"},
          "e2e": {"value": n / s_e2e, "unit": "samples/s", "h2d_bytes_per_step": sum(t.numel() * 4 for t in host),
                  "d2h_bytes_per_step": 4 * n},
          "samples_axis": {"value": n / s_k, "unit": "samples/s", "joint_ll": ll_k,
                           "what": "the same pass sharded over the K importance samples (every rank sees all datapoints, "
                                   "K/N samples each), combined with one MAX + one SUM all-reduce per evaluator"},
          "achieved_tflops_algorithmic": n * EVAL_FLOP_PER_DATAPOINT / s / 1e12}
    # per-kernel accounting of one evaluator pass on a slice (the pass is a loop over datapoint chunks of identical launches)
    m = min(n, 32)
    small = [t[:m] for t in data]
    agg, total, _ = profile_pass(lambda: [P.conditional_likelihood(model, small, 0, 1, K=1000, batch_size_K=100, axis="batch"),
                                          P.conditional_likelihood(model, small, 1, 0, K=1000, batch_size_K=100, axis="batch"),
                                          P.joint_likelihood(model, small, K=1000, batch_size_K=100, axis="batch")])
    roofs = roofline_of(agg, total, peak)
    ev["roofline"] = roofs[0] if roofs else None
    ev["rooflines"] = roofs
    ev["kernel_shares"] = {k: {"ms": round(a["ms"], 4), "launches": a["n"], "share": round(a["ms"] / total, 4)}
                           for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"])[:12]}
    return ev


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
