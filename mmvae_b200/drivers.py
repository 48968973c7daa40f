"""Training and evaluation drivers: the callers of the hot path, torchrun-able.

    python -m mmvae_b200.drivers train --config-path cfg.json [--synthetic 6000] [--out runs/x]
    python -m mmvae_b200.drivers eval  --run runs/x [--k 1000] [--repeats 5]
    torchrun --nproc-per-node N -m mmvae_b200.drivers train ...          (batch-sharded, one all-reduce per step)

`train` is main.py:24-279 reduced to what touches the path: JSON config -> `VAE_<model>` -> Adam(amsgrad, lr) ->
epochs of objective/backward/step (through train.Trainer: flat buffers, CUDA-graph replay for full batches) ->
validation objective under no_grad -> ReduceLROnPlateau('min') -> best-validation checkpoint `model.pt` (reference
state_dict keys) + `args.json` + `losses.pt` -> early stop after 20 epochs without improvement (main.py:234-277).
`eval` is compute_likelihoods.py:82-122: loader batch 500, conditional + joint K-sample likelihoods per batch, mean and
std over batches (`get_mean_std`, utils.py:253-255), repeated `--repeats` times.
wandb, plots, samplers and the analysis hooks are out of scope.
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

from . import data as D
from . import objectives
from . import parallel as P


# ----------------------------------------------------------------------------------------------------- helpers
def load_config(path):
    """main.py:33-36: the JSON file is the argparse Namespace."""
    with open(path) as f:
        args = argparse.Namespace()
        args.__dict__.update(json.load(f))
    return apply_defaults(args)


def apply_defaults(args):
    for k, v in (("seed", 1), ("K", 1), ("looser", False), ("obj", "elbo"), ("learning_rate", 1e-3), ("epochs", 1),
                 ("batch_size", 128), ("warmup", 0), ("beta_prior", 1), ("llik_scaling", 0.0), ("dist", "normal"),
                 ("recon_losses", None), ("data_path", ""), ("latent_dim", 20)):
        if not hasattr(args, k):
            setattr(args, k, v)
    return args


def objective_name(args, multimodal=True):
    """main.py:134-137."""
    return ("m_" if multimodal else "") + args.obj + ("_looser" if (args.looser and args.obj != "elbo") else "")


def update_details(agg, new):
    """utils.py `update_details`: running sums of the per-batch detail scalars."""
    for k, v in new.items():
        v = float(v.sum()) if torch.is_tensor(v) else float(v)
        agg[k] = agg.get(k, 0.0) + v


def get_mean_std(d):
    """utils.py:253-255."""
    return ({k: torch.mean(torch.tensor(v)) for k, v in d.items()},
            {"std_" + k: torch.std(torch.tensor(v)) for k, v in d.items()})


class ReduceLROnPlateau:
    """torch.optim.lr_scheduler.ReduceLROnPlateau('min') defaults (factor 0.1, patience 10, rel threshold 1e-4), acting
    on the trainer's learning rate (main.py:117,258)."""

    def __init__(self, trainer, factor=0.1, patience=10, threshold=1e-4, min_lr=0.0):
        self.t, self.factor, self.patience, self.threshold, self.min_lr = trainer, factor, patience, threshold, min_lr
        self.best, self.bad = float("inf"), 0

    def step(self, metric):
        if metric < self.best * (1.0 - self.threshold):
            self.best, self.bad = metric, 0
        else:
            self.bad += 1
        if self.bad > self.patience:
            self.t.set_lr(max(self.t.lr * self.factor, self.min_lr))
            self.bad = 0


def _dist_setup():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, torch.device("cuda", local)


def build_model(args, device):
    import mmvae_b200
    apply_defaults(args)
    if args.recon_losses is None:
        args.recon_losses = ["normal"] * len(getattr(mmvae_b200, "VAE_" + args.model).shape_mods)
    torch.manual_seed(args.seed)                                                  # main.py:49-51
    return getattr(mmvae_b200, "VAE_" + args.model)(args).to(device)              # main.py:70-71


def build_loaders(args, model, batch_size, device, rank, world, synthetic=None):
    """train / test / val batches.  With `synthetic` = N the base datasets are stand-ins of the model's shapes with N
    training images per modality; otherwise `args.data_path` must hold `<name>_{train,test}.pt` image tensors, their
    `<name>_{train,test}_labels.pt`, and (optionally) the pair-index files of data.make_pair_indices."""
    shapes = model.shape_mods
    if synthetic:
        train_b, test_b = D.synthetic_bases(shapes, n_train=synthetic, n_test=max(500, synthetic // 6), seed=args.seed)
        g = torch.Generator().manual_seed(args.seed)
        tr_idx = D.rand_match_on_idx([l for _, l in train_b], dm=5, generator=g)
        te_idx = D.rand_match_on_idx([l for _, l in test_b], dm=5, generator=g)
    else:
        names = [v.modelName for v in model.vaes]
        ld = lambda n, s: torch.load(os.path.join(args.data_path, f"{n}_{s}.pt"))              # noqa: E731
        train_b = [(ld(n, "train"), ld(n, "train_labels")) for n in names]
        test_b = [(ld(n, "test"), ld(n, "test_labels")) for n in names]
        try:
            tr_idx, te_idx = D.load_pair_indices(args.data_path, len(names))
        except RuntimeError:
            tr_idx, te_idx = D.make_pair_indices(args.data_path, [l for _, l in train_b], [l for _, l in test_b],
                                                 seed=args.seed)
    return D.mnist_svhn_like_loaders(train_b, test_b, tr_idx, te_idx, batch_size, True, device,
                                     getattr(args, "len_train", None), rank, world)


# ------------------------------------------------------------------------------------------------------- train
def run_training(args, out_dir, synthetic=None, max_steps=None, log=print):
    from .train import Trainer
    rank, world, device = _dist_setup()
    model = build_model(args, device)
    obj_name = objective_name(args, hasattr(model, "vaes"))
    objective = getattr(objectives, obj_name)
    per_rank = max(1, args.batch_size // world)              # the reference config's GLOBAL batch, sharded (strong)
    train_loader, test_loader, val_loader = build_loaders(args, model, per_rank, device, rank, world, synthetic)
    trainer = Trainer(model, obj_name, K=args.K, lr=args.learning_rate, amsgrad=True)
    trainer.broadcast_parameters(0)
    scheduler = ReduceLROnPlateau(trainer)
    if rank == 0:
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "args.json"), "w") as f:
            json.dump({k: v for k, v in args.__dict__.items() if isinstance(v, (int, float, str, bool, list, type(None)))}, f)
    log(f"model {args.model}  objective {obj_name}  K={args.K}  world={world}  batch/rank={per_rank}  "
        f"train pairs {train_loader.dataset_len}  val {val_loader.dataset_len}  test {test_loader.dataset_len}")
    agg = {"train_loss": [], "test_loss": [], "samples_per_s": []}
    best, bad, steps = float("inf"), 0, 0
    for epoch in range(1, args.epochs + 1):
        model.train()
        trainer.set_schedule(epoch, args.warmup, args.beta_prior)       # main.py:155: objective(model, data, K, epoch, warmup, beta_prior)
        b_loss = torch.zeros((), device=device)
        n_seen, t0 = 0, time.perf_counter()
        for x in train_loader:
            full = x[0].shape[0] == per_rank
            if full and trainer.graph_ready(x):
                loss = trainer.step_graph(x)                 # CUDA-graph replay (static shapes)
            else:
                loss = trainer.step(x)                       # ragged tail batch / first step: eager launches
                if full and not trainer.graph_ready(x):
                    trainer.capture(x, warmup=1)
            b_loss -= loss                                   # main.py:156-160: minimised loss = -objective
            n_seen += x[0].shape[0]
            steps += 1
            if max_steps and steps >= max_steps:
                break
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if hasattr(model, "step"):
            model.step(epoch)                                # main.py:171: the model's own per-epoch schedule (beta_kl decay, ...)
        tot = torch.stack([b_loss, torch.tensor(float(n_seen), device=device)])
        P.allreduce_sum_(tot)
        train_loss = (tot[0] / tot[1]).item()
        agg["train_loss"].append(train_loss)
        agg["samples_per_s"].append(tot[1].item() / dt)
        # validation (main.py:174-215): the same objective under no_grad
        model.eval()
        v = torch.zeros(2, device=device)
        with torch.no_grad():
            for x in val_loader:
                l, _ = objective(model, x, args.K, epoch, args.warmup, args.beta_prior)
                v[0] -= l
                v[1] += x[0].shape[0]
        P.allreduce_sum_(v)
        val_loss = (v[0] / v[1].clamp_min(1)).item()
        agg["test_loss"].append(val_loss)
        log(f"====> Epoch: {epoch:03d} Train loss: {train_loss:.4f}  Test loss: {val_loss:.4f}  "
            f"{agg['samples_per_s'][-1]:.0f} samples/s  lr {trainer.lr:g}")
        if val_loss < best:                                  # main.py:240-246
            bad, best = 0, val_loss
            if rank == 0:
                torch.save(model.state_dict(), os.path.join(out_dir, "model.pt"))
        else:
            bad += 1
        scheduler.step(val_loss)
        if rank == 0:
            torch.save(agg, os.path.join(out_dir, "losses.pt"))
        if bad == 20 and epoch >= args.warmup:               # main.py:260-263
            break
        if max_steps and steps >= max_steps:
            break
    return agg


# -------------------------------------------------------------------------------------------------------- eval
def run_likelihoods(args, run_dir, K=1000, repeats=5, synthetic=None, batch_size=500, batch_size_K=100,
                    max_batches=None, log=print):
    """compute_likelihoods.py:82-122.  Multi-rank: every batch is sharded over ranks by datapoints (parallel.py)."""
    rank, world, device = _dist_setup()
    model = build_model(args, device)
    sd_path = os.path.join(run_dir, "model.pt")
    if os.path.exists(sd_path):
        model.load_state_dict(torch.load(sd_path, map_location=device))
    model.eval()
    _, test_loader, _ = build_loaders(args, model, batch_size, device, 0, 1, synthetic)   # same batches on every rank
    results = []
    for r in range(repeats):
        b_metrics = {}
        t0, n = time.perf_counter(), 0
        for i, x in enumerate(test_loader):
            if max_batches and i >= max_batches:
                break
            m, sd = {}, 1 + r * 100003 + i                     # fresh, rank-independent draws per run and batch
            for c in range(model.mod):
                for g in range(model.mod):
                    if c != g:
                        m.update(P.conditional_likelihood(model, x, c, g, K=K, batch_size_K=batch_size_K, axis="batch",
                                                          seed=sd)[0])
            if model.mod == 2:                                   # the reference's mixture sampling is bimodal (mmvae.py:128)
                m.update(P.joint_likelihood(model, x, K=K, batch_size_K=batch_size_K, axis="batch", seed=sd)[0])
            for k, v in m.items():
                b_metrics.setdefault(k, []).append(float(v))
            n += x[0].shape[0]
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        mean, std = get_mean_std(b_metrics)
        results.append({"mean": {k: float(v) for k, v in mean.items()}, "std": {k: float(v) for k, v in std.items()},
                        "datapoints_per_s": n / dt})
        log(f"run {r}: " + "  ".join(f"{k} {float(v):.3f} +- {float(std['std_' + k]):.3f}" for k, v in mean.items())
            + f"   [{n / dt:.1f} datapoints/s]")
    if rank == 0:
        with open(os.path.join(run_dir, f"likelihoods_K{K}.json"), "w") as f:
            json.dump(results, f)
    return results


def main(argv=None):
    ap = argparse.ArgumentParser(description="mmvae_b200 drivers (main.py / compute_likelihoods.py of the reference)")
    sub = ap.add_subparsers(dest="cmd", required=True)
    t = sub.add_parser("train")
    t.add_argument("--config-path", required=True)
    t.add_argument("--out", default="runs/last")
    t.add_argument("--synthetic", type=int, default=0, help="N synthetic training images per modality (no datasets)")
    t.add_argument("--max-steps", type=int, default=0)
    e = sub.add_parser("eval")
    e.add_argument("--run", required=True)
    e.add_argument("--k", type=int, default=1000)
    e.add_argument("--repeats", type=int, default=5)
    e.add_argument("--synthetic", type=int, default=0)
    e.add_argument("--max-batches", type=int, default=0)
    a = ap.parse_args(argv)
    rank = int(os.environ.get("RANK", "0"))
    log = print if rank == 0 else (lambda *_: None)
    if a.cmd == "train":
        run_training(load_config(a.config_path), a.out, a.synthetic or None, a.max_steps or None, log)
    else:
        run_likelihoods(load_config(os.path.join(a.run, "args.json")), a.run, a.k, a.repeats, a.synthetic or None,
                        max_batches=a.max_batches or None, log=log)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
