"""Training step of the hot path (main.py:114-115,147-171): objective -> backward -> [gradient all-reduce] ->
Adam(amsgrad).  Parameters and gradients live in two flat fp32 buffers so the optimizer is ONE fused kernel and the
multi-GPU exchange is ONE NCCL all-reduce (SUM — the reference objective is an un-normalised sum over the batch,
objectives.py:403,438, so batch shards add).  The whole step can be captured in a CUDA graph.
"""
import torch
import torch.distributed as dist

from . import objectives, ops


class FlatParams:
    """Re-homes every trainable parameter of `model` (and its .grad) as views into two flat buffers."""

    def __init__(self, model):
        self.params = [p for p in model.parameters() if p.requires_grad]
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self.flat = torch.empty(n, device=dev, dtype=torch.float32)
        self.grad = torch.zeros(n, device=dev, dtype=torch.float32)
        o = 0
        for p in self.params:
            k = p.numel()
            self.flat[o:o + k].copy_(p.data.reshape(-1))
            p.data = self.flat[o:o + k].view_as(p.data)
            p.grad = self.grad[o:o + k].view_as(p.data)
            o += k
        self.numel = n


class Trainer:
    """One data-parallel training step.  `objective` is a name from mmvae_b200.objectives ('m_dreg_looser', ...)."""

    def __init__(self, model, objective="m_dreg_looser", K=30, lr=1e-3, amsgrad=True, process_group=None,
                 distributed=None, branch_streams=True, grad_sink=True, wgrad_stream=True):
        self.model = model
        self.objective = getattr(objectives, objective)
        self.K = K
        self.lr, self.amsgrad = lr, amsgrad
        self.fp = FlatParams(model)
        n, dev = self.fp.numel, self.fp.flat.device
        self.m = torch.zeros(n, device=dev)
        self.v = torch.zeros(n, device=dev)
        self.vmax = torch.zeros(n, device=dev)
        self.step_dev = torch.zeros(1, device=dev)
        self.pg = process_group
        self.distributed = dist.is_initialized() and dist.get_world_size() > 1 if distributed is None else distributed
        self._graph = None
        # objective(model, x, K, epoch, warmup, beta_prior) as main.py:155 calls it; set_schedule() updates these per epoch
        self.epoch, self.warmup, self.beta_prior = 1, 0, 1
        self._sched_key = None
        self.grad_sink = grad_sink
        self._wgrad_stream = torch.cuda.Stream(device=dev) if (wgrad_stream and grad_sink) else None
        if branch_streams and hasattr(model, "_run_branches"):
            model.branch_streams = True                      # per-modality chains on side streams (models.py)

    def broadcast_parameters(self, src=0):
        if self.distributed:
            dist.broadcast(self.fp.flat, src, group=self.pg)

    def set_schedule(self, epoch, warmup=0, beta_prior=1):
        """The positional arguments the reference passes to every objective (main.py:155: epoch, args.warmup,
        args.beta_prior).  m_jmvae_nf turns them into its two training phases, the frozen sub-modules of the second one and
        the beta schedules, all of which a captured graph bakes in (requires_grad pattern, Python scalars): when anything the
        objective can see changes, the graph is dropped and the caller re-captures on its next full batch."""
        self.epoch, self.warmup, self.beta_prior = epoch, warmup, beta_prior
        key = (epoch >= warmup, epoch if getattr(self.model, "linear_warmup", False) else None, beta_prior,
               getattr(self.model, "beta_kl", None))
        if key != self._sched_key:
            if self._sched_key is not None:
                self._graph = None
                self._graph_upd = None
            self._sched_key = key

    def _fwd_bwd(self, x, noise=None):
        self.fp.grad.zero_()
        loss, _ = self.objective(self.model, x, self.K, self.epoch, self.warmup, self.beta_prior, noise=noise)
        ops.GRAD_SINK = set() if self.grad_sink else None    # weight/bias gradients land in the flat buffer directly
        ops.WGRAD_STREAM = self._wgrad_stream if self.grad_sink else None   # ... computed off the data-gradient chain
        try:
            (-loss).backward()                               # main.py:156: the objective is maximised
        finally:
            ops.GRAD_SINK = None
            ops.WGRAD_STREAM = None
        if hasattr(self.model, "join_branch_streams"):
            self.model.join_branch_streams()
        if ops.WGRAD_USED:
            torch.cuda.current_stream().wait_stream(self._wgrad_stream)
            ops.WGRAD_USED = False
        return loss.detach()

    def _reduce(self):
        if self.distributed:                                 # ONE collective per step: the flat gradient bucket
            dist.all_reduce(self.fp.grad, op=dist.ReduceOp.SUM, group=self.pg)

    def _active_segments(self):
        """Contiguous ranges of the flat buffer whose parameters currently require a gradient.  torch.optim.Adam skips a
        parameter whose .grad is None (the frozen joint encoder / decoders of m_jmvae_nf's second phase, objectives.py:180-183):
        its moments and the parameter stay put; updating it with a zero gradient would keep moving it along its momentum."""
        flags = tuple(p.requires_grad for p in self.fp.params)
        if getattr(self, "_seg_flags", None) != flags:
            segs, o, start = [], 0, None
            for p, f in zip(self.fp.params, flags):
                if f and start is None:
                    start = o
                if not f and start is not None:
                    segs.append((start, o))
                    start = None
                o += p.numel()
            if start is not None:
                segs.append((start, o))
            self._seg_flags, self._segs = flags, segs
        return self._segs

    def _update(self):
        segs = self._active_segments()
        for i, (a, b) in enumerate(segs):
            step = self.step_dev
            if i > 0:                                        # the kernel increments the step counter it is given: once per step
                step = self.step_dev - 1.0
            ops.adam_amsgrad_(self.fp.flat[a:b], self.fp.grad[a:b], self.m[a:b], self.v[a:b], self.vmax[a:b], step, lr=self.lr,
                              amsgrad=self.amsgrad)

    def _step_impl(self, x, noise=None):
        loss = self._fwd_bwd(x, noise)
        self._reduce()
        self._update()
        return loss

    def step(self, x, noise=None):
        """eager step -> loss (device scalar, this rank's shard)."""
        return self._step_impl(x, noise)

    # ---- CUDA-graph replay: static input buffers, RNG inside the graph -------------------------------------------
    # Two graphs (forward+backward, optimizer) with the NCCL all-reduce issued between them on the same stream: the collective
    # stays an ordinary stream-ordered launch, the ~100 kernel launches around it are replayed.  capture_collective=True puts
    # the all-reduce INSIDE one graph (no host gap around it); measured once at N=2 (2.231 ms/step) but two of three further
    # multi-rank runs with it hung on this stack (the communicator's watchdog thread and global-mode capture), so it is opt-in.
    def capture(self, x_example, warmup=3, capture_collective=False):
        self._static_x = [t.clone() for t in x_example]
        # the warm-up iterations run real steps (allocator and lazy initialisations must happen before capture): parameters,
        # moments and the step counter are put back afterwards, so capturing is invisible to the optimisation trajectory
        saved = [t.clone() for t in (self.fp.flat, self.m, self.v, self.vmax, self.step_dev)]
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self._step_impl(self._static_x)
        torch.cuda.current_stream().wait_stream(s)
        for dst, src in zip((self.fp.flat, self.m, self.v, self.vmax, self.step_dev), saved):
            dst.copy_(src)
        ops.weights_changed()                                # derived weight layouts must be rebuilt INSIDE the graph
        one_graph = capture_collective or not self.distributed
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._static_loss = self._fwd_bwd(self._static_x)
            if one_graph:
                self._reduce()
                self._update()
        self._graph_upd = None
        if not one_graph:
            self._graph_upd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph_upd):
                self._update()
        return self

    def graph_ready(self, x):
        """True when a captured graph exists for batches shaped like x."""
        return self._graph is not None and all(a.shape == b.shape for a, b in zip(self._static_x, x))

    def set_lr(self, lr):
        """The learning rate is a launch argument of the fused optimizer kernel, i.e. baked into a captured graph: a
        change (ReduceLROnPlateau, main.py:258) drops the graph; the caller re-captures on its next full batch."""
        if lr != self.lr:
            self.lr = lr
            self._graph = None
            self._graph_upd = None

    def step_graph(self, x):
        for dst, src in zip(self._static_x, x):
            dst.copy_(src, non_blocking=True)
        ops.weights_changed()                                # the replayed optimizer writes the parameters in place
        self._graph.replay()
        if self._graph_upd is not None:
            self._reduce()
            self._graph_upd.replay()
        return self._static_loss
