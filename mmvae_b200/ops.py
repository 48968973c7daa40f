"""torch.autograd bindings of the C-ABI kernels (forward + registered backward).  Device fp32 only; every op raises
if handed a CPU tensor — there is no eager / CPU fallback on the product path.

Layouts: see include/mmvae_b200.h.  Image activations are carried as tensors of LOGICAL shape (N,C,H,W) whose memory
is channels-last (NHWC) between layers; the lowering kernels take explicit strides so NCHW tensors at the boundary
need no transpose pass.
"""
import ctypes
import math
import weakref

import torch

from . import _lib

C = _lib.CONSTS
DIST = {"normal": C["MMVAE_DIST_NORMAL"], "laplace": C["MMVAE_DIST_LAPLACE"]}
LIK = {"normal": C["MMVAE_LIK_NORMAL"], "laplace": C["MMVAE_LIK_LAPLACE"], "bernoulli": C["MMVAE_LIK_BERNOULLI"],
       "neg_mse": C["MMVAE_LIK_NEG_MSE"], "neg_l1": C["MMVAE_LIK_NEG_L1"], "neg_bce": C["MMVAE_LIK_NEG_BCE"]}
ACT = {None: C["MMVAE_ACT_NONE"], "none": C["MMVAE_ACT_NONE"], "relu": C["MMVAE_ACT_RELU"],
       "sigmoid": C["MMVAE_ACT_SIGMOID"]}
XFORM = {"exp_half": C["MMVAE_XFORM_EXP_HALF"], "softmax": C["MMVAE_XFORM_SOFTMAX"],
         "exp_half_eps": C["MMVAE_XFORM_EXP_HALF_EPS"]}
OBJ = {"iwae": C["MMVAE_OBJ_IWAE"], "iwae_looser": C["MMVAE_OBJ_IWAE_LOOSER"], "dreg": C["MMVAE_OBJ_DREG"],
       "dreg_looser": C["MMVAE_OBJ_DREG_LOOSER"]}

# launch counter (bench.py reports gpu_launches from it); counts C-ABI calls, each >= 1 kernel launch
N_CALLS = 0
ALLOW_TC = True          # let mmvae_gemm pick the tcgen05 path where it supports the shape
ALLOW_TS = True          # transposed convolutions: second-generation kernels (A operand in tensor memory, bf16x3 split)


def _dev(t, name="tensor"):
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32):
        raise RuntimeError(f"mmvae_b200: {name} must be a CUDA float32 tensor (no CPU fallback); got "
                           f"{type(t).__name__} {getattr(t, 'device', None)} {getattr(t, 'dtype', None)}")
    return t


def _c(t, name="tensor"):
    _dev(t, name)
    return t if t.is_contiguous() else t.contiguous()


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


GRAD_SINK = None         # a set() while train.Trainer runs backward: see _sink()


def _sink(p):
    """Destination for the gradient of parameter `p`, or None.  While GRAD_SINK is a set, the FIRST gradient of a leaf
    parameter in a backward pass is written by the producing kernel straight into the (zeroed) `p.grad` buffer and the
    autograd Function returns None for it, so no AccumulateGrad `+=` kernel runs; later uses of the same parameter in
    the same pass return their gradient normally and autograd adds them."""
    if GRAD_SINK is None or p is None or not p.is_leaf or p.grad is None or not p.grad.is_contiguous():
        return None
    key = p.data_ptr()
    if key in GRAD_SINK:
        # second use of the parameter in this backward pass: autograd adds this gradient to p.grad on the current stream,
        # while the first one may still be being WRITTEN by a kernel on the weight-gradient stream -- order them
        if WGRAD_STREAM is not None and WGRAD_USED:
            torch.cuda.current_stream().wait_stream(WGRAD_STREAM)
        return None
    GRAD_SINK.add(key)
    return p.grad


WGRAD_STREAM = None      # torch.cuda.Stream while train.Trainer runs backward: see _wgrad_side()
WGRAD_USED = False       # set when a kernel was issued on WGRAD_STREAM (the trainer joins the stream only then)


class _wgrad_side:
    """Context for the weight / bias gradient kernels of one backward node.  They do not feed the rest of the backward
    pass (only the optimizer reads them), so when their results are sunk into the flat gradient buffer (`sunk`) and a
    WGRAD_STREAM is set they are issued on that stream: the data-gradient chain continues on the node's own stream and
    the latency-bound launches further down (encoders) hide under the large weight-gradient kernels.  `reads` are the
    tensors those kernels read (kept alive for the side stream); the trainer joins the stream before the optimizer."""

    def __init__(self, sunk, *reads):
        self.on = bool(sunk) and WGRAD_STREAM is not None
        self.reads = reads

    def __enter__(self):
        if self.on:
            global WGRAD_USED
            WGRAD_USED = True
            s = WGRAD_STREAM
            s.wait_stream(torch.cuda.current_stream())
            for t in self.reads:
                if t is not None:
                    t.record_stream(s)
            self.ctx = torch.cuda.stream(s)
            self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.on:
            self.ctx.__exit__(*exc)
        return False


def _sinks(ctx_leaves, need_w, need_b):
    """-> (dst_w, dst_b, all_sunk): sink buffers of the weight / bias gradients of a node (None = return it to autograd)."""
    dw = _sink(ctx_leaves[0]) if need_w else None
    db = _sink(ctx_leaves[1]) if need_b else None
    return dw, db, (dw is not None or not need_w) and (db is not None or not need_b) and (need_w or need_b)


RELU_TAP = None          # test hook: set to a dict to record, per weight (data_ptr), the sign pattern y > 0 of every fused-ReLU
                         # output in call order -- lets a double-precision checker be evaluated on the SAME side of each kink (tests only)


def _tap_relu(w, b, y, act):
    if RELU_TAP is not None and act == "relu":
        m = y.detach() > 0
        RELU_TAP.setdefault(w.data_ptr(), []).append(m)
        if b is not None:                       # masked (MADE) weights are temporaries: the bias identifies the layer
            RELU_TAP.setdefault(b.data_ptr(), []).append(m)


GEMM_KERNELS = {0: "simt", 1: "tcgen05", 2: "skinny", 3: "small"}
GEMM_LOG = None          # set to a list (with PROFILE) to record ("M N K a_kmajor b_kmajor act", start, end) per GEMM
PROFILE = None           # set to a list to record (name, work, start_event, end_event) per C-ABI call (bench.py)


def _call(name, *args, work=None):
    """work: ('flop'|'byte', amount) — the ALGORITHMIC work of this call, used for roofline accounting."""
    global N_CALLS
    N_CALLS += 1
    if PROFILE is None:
        _lib.check(getattr(_lib.load(), name)(*args), name)
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(getattr(_lib.load(), name)(*args), name)
    e1.record()
    PROFILE.append((name, work, e0, e1))


# ------------------------------------------------------------------------------------------------ raw wrappers
def gemm(a, b, bias, M, N, K, a_kmajor, b_kmajor, act=None, out=None):
    """C[M,N] = act(A·B + bias); a/b are 2-D contiguous device tensors in the stated storage order."""
    lib = _lib.load()
    out = torch.empty((M, N), device=a.device, dtype=torch.float32) if out is None else out
    ws_bytes = lib.mmvae_gemm_workspace_bytes(M, N, K, int(a_kmajor), int(b_kmajor)) if ALLOW_TC else \
        lib.mmvae_gemm_workspace_bytes(M, N, K, -1, -1)
    ws = torch.empty(ws_bytes // 4, device=a.device, dtype=torch.float32) if ws_bytes else None
    _call("mmvae_gemm", _p(a), _p(b), _p(bias), _p(out), M, N, K, int(a_kmajor), int(b_kmajor), ACT[act], _p(ws),
          ws_bytes, int(ALLOW_TC), _stream(), work=("flop", 2.0 * M * N * K))
    if PROFILE:                                   # which kernel the dispatcher picked: the per-kernel accounting of bench.py
        PROFILE[-1] = ("mmvae_gemm[%s]" % GEMM_KERNELS[lib.mmvae_gemm_last_kernel()],) + PROFILE[-1][1:]
    if GEMM_LOG is not None and PROFILE:
        GEMM_LOG.append((f"M={M} N={N} K={K} a_k={int(a_kmajor)} b_k={int(b_kmajor)} bias={bias is not None} act={act}",
                         PROFILE[-1][2], PROFILE[-1][3]))
    return out


def colsum(a2d, out=None):
    M, N = a2d.shape
    lib = _lib.load()
    ws = torch.empty(lib.mmvae_colsum_workspace_bytes(M, N) // 4, device=a2d.device, dtype=torch.float32)
    out = torch.empty(N, device=a2d.device, dtype=torch.float32) if out is None else out
    _call("mmvae_colsum", _p(a2d), _p(out), _p(ws), M, N, _stream())
    return out


def convt_fwd_loglik_supported(w, H, W, lik):
    Cin, Cout = w.shape[0], w.shape[1]
    return w.shape[2] == 4 and w.shape[3] == 4 and lik in LIK and \
        _lib.load().mmvae_convt_small_fwd_loglik_supported(H, W, Cin, Cout, LIK[lik]) == 1


def convt_small_fwd_loglik(x, w, b, act, x_flat, x_div, lik, scale):
    """-> lp [N]: the image-producing ConvTranspose2d + activation fused with sum_pix log p(x_flat[(n / x_div) % B] | y_n)."""
    N, Cin, H, W = x.shape
    Cout = w.shape[1]
    x2d = _as_rows(_dev(x, "x"))
    xf = _c(x_flat, "x")
    lp = torch.empty(N, device=x.device, dtype=torch.float32)
    _call("mmvae_convt_small_fwd_loglik", _p(x2d), _p(_c(w, "weight")), _p(b), _p(xf), _p(lp), None, N, H, W, Cin, Cout, ACT[act],
          xf.shape[0], int(x_div), LIK[lik], float(scale), _stream(), work=("byte", 4.0 * x2d.numel()))
    return lp


def act_bwd(g, y, act):
    if ACT[act] == C["MMVAE_ACT_NONE"]:
        return g
    out = torch.empty_like(g)
    _call("mmvae_act_bwd", _p(g), _p(y), _p(out), g.numel(), ACT[act], _stream(), work=("byte", 12.0 * g.numel()))
    return out


# Derived weight layouts (lowered / class-major / bf16-packed copies of a conv weight) are pure functions of the weight.  The
# K=1000 evaluators call each decoder ~10 times per batch with frozen weights (12 permute + 4 pack launches per pass, 6 % of
# it): under torch.no_grad() a layout is built once per weight VALUE and reused while the SAME tensor object is unchanged --
# tracked by a weak reference to the object (a new model at a recycled address is a different object), its version counter
# (any torch in-place update) and WEIGHT_EPOCH (bumped by the fused optimizer and by graph replays, which write through raw
# pointers).  With gradients enabled nothing is cached: a training step uses each layout once.
WEIGHT_EPOCH = 0
_LAYOUTS = {}


def weights_changed():
    """Call after writing parameters behind torch's back (fused optimizer kernel, CUDA-graph replay)."""
    global WEIGHT_EPOCH
    WEIGHT_EPOCH += 1


def _layout(w, kind, build):
    if torch.is_grad_enabled() or torch.cuda.is_current_stream_capturing():
        return build()
    key = (id(w), kind)
    stamp = (w._version, WEIGHT_EPOCH, w.data_ptr(), torch.cuda.current_stream().cuda_stream)
    hit = _LAYOUTS.get(key)
    if hit is not None and hit[0]() is w and hit[1] == stamp:
        return hit[2]
    out = build()
    if len(_LAYOUTS) > 64:
        for k in [k for k, v in _LAYOUTS.items() if v[0]() is None]:
            del _LAYOUTS[k]
    _LAYOUTS[key] = (weakref.ref(w), stamp, out)
    return out


def permute_weight(w):
    """torch conv weight [C0,C1,kh,kw] -> lowered 2-D [C0, kh*kw*C1] (tap-major, C1 fastest)."""
    def build():
        C0, C1, kh, kw = w.shape
        out = torch.empty((C0, kh * kw * C1), device=w.device, dtype=torch.float32)
        _call("mmvae_permute_weight", _p(w), _p(out), C0, C1, kh * kw, 1, _stream())
        return out
    return _layout(w, "permuted", build)


def unpermute_weight(wp, shape, out=None):
    C0, C1, kh, kw = shape
    out = torch.empty(shape, device=wp.device, dtype=torch.float32) if out is None else out
    _call("mmvae_permute_weight", _p(wp), _p(out), C0, C1, kh * kw, 0, _stream())
    return out


def _strides_nchw(t):
    """element strides (sn, sc, sy, sx) of a logical (N,C,H,W) tensor."""
    return t.stride(0), t.stride(1), t.stride(2), t.stride(3)


def _dense_image(t):
    """logical (N,C,H,W) tensor usable by the lowering kernels: non-overlapping and dense in some order."""
    _dev(t, "image")
    if t.is_contiguous() or t.is_contiguous(memory_format=torch.channels_last):
        return t
    if t.permute(0, 2, 3, 1).is_contiguous():
        return t
    return t.contiguous()


def im2col(x, k, s, p):
    N, Cc, H, W = x.shape
    Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    col = torch.empty((N * Ho * Wo, k * k * Cc), device=x.device, dtype=torch.float32)
    sn, sc, sy, sx = _strides_nchw(x)
    _call("mmvae_im2col", _p(x), _p(col), N, Cc, H, W, k, s, p, sn, sc, sy, sx, _stream(),
          work=("byte", 4.0 * (x.numel() + col.numel())))
    return col, Ho, Wo


def col2im(col, bias, N, Cc, H, W, k, s, p, act=None, nchw_out=False):
    """col [N*H*W, k*k*C] -> logical (N,C,Ho,Wo); memory NHWC unless nchw_out."""
    Ho, Wo = (H - 1) * s - 2 * p + k, (W - 1) * s - 2 * p + k
    if nchw_out:
        y = torch.empty((N, Cc, Ho, Wo), device=col.device, dtype=torch.float32)
    else:
        y = torch.empty((N, Ho, Wo, Cc), device=col.device, dtype=torch.float32).permute(0, 3, 1, 2)
    sn, sc, sy, sx = _strides_nchw(y)
    _call("mmvae_col2im", _p(col), _p(bias), _p(y), N, Cc, H, W, k, s, p, sn, sc, sy, sx, ACT[act], _stream(),
          work=("byte", 4.0 * (col.numel() + y.numel())))
    return y


def _as_rows(t):
    """logical (N,C,H,W) -> contiguous matrix [N*H*W, C] (free when the memory is already NHWC)."""
    N, Cc, H, W = t.shape
    return t.permute(0, 2, 3, 1).contiguous().view(N * H * W, Cc)


# ------------------------------------------------------------------------------------------- autograd functions
class LinearAct(torch.autograd.Function):
    """y = act(x @ W^T + b)   (torch.nn.Linear + ReLU/Sigmoid; pythae MLPs, joint_encoders.py:61-82)."""

    @staticmethod
    def forward(ctx, x, w, b, act):
        x, w = _c(x, "x"), _c(w, "weight")
        R, K = x.shape
        N = w.shape[0]
        y = gemm(x, w, b, R, N, K, True, True, act)
        ctx.act = act
        ctx.save_for_backward(x, w, y if ACT[act] != C["MMVAE_ACT_NONE"] else None)
        ctx.has_bias = b is not None
        ctx.leaves = (w, b)
        _tap_relu(w, b, y, act)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w, y = ctx.saved_tensors
        g = act_bwd(_c(gy, "grad"), y, ctx.act)
        R, K = x.shape
        N = w.shape[0]
        gx = gemm(g, w, None, R, K, N, True, False) if ctx.needs_input_grad[0] else None
        gw = gb = None
        need_w, need_b = ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2]
        dw, db, sunk = _sinks(ctx.leaves, need_w, need_b)
        with _wgrad_side(sunk, g, x):
            if need_w:
                gw = gemm(g, x, None, N, K, R, False, False, out=dw)
                gw = None if dw is not None else gw
            if need_b:
                gb = colsum(g, out=db)
                gb = None if db is not None else gb
        return gx, gw, gb, None


class LinearPreact(torch.autograd.Function):
    """Final decoder layer whose Sigmoid backward is fused into the log-likelihood backward: forward is
    y = sigmoid(x W^T + b); backward receives the gradient w.r.t. the PRE-activation (see LogLik through_sigmoid)."""

    @staticmethod
    def forward(ctx, x, w, b):
        x, w = _c(x, "x"), _c(w, "weight")
        y = gemm(x, w, b, x.shape[0], w.shape[0], x.shape[1], True, True, "sigmoid")
        ctx.save_for_backward(x, w)
        ctx.leaves = (w, b)
        return y

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        g = _c(g, "grad")
        R, K = x.shape
        N = w.shape[0]
        gx = gemm(g, w, None, R, K, N, True, False) if ctx.needs_input_grad[0] else None
        gw = gb = None
        need_w, need_b = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        dw, db, sunk = _sinks(ctx.leaves, need_w, need_b)
        with _wgrad_side(sunk, g, x):
            if need_w:
                gw = gemm(g, x, None, N, K, R, False, False, out=dw)
                gw = None if dw is not None else gw
            if need_b:
                gb = colsum(g, out=db)
                gb = None if db is not None else gb
        return gx, gw, gb


class Conv2dAct(torch.autograd.Function):
    """y = act(conv2d(x, W, b, stride, pad))  (encoders.py:80-105).  Output memory NHWC, logical (N,Cout,Ho,Wo)."""

    @staticmethod
    def forward(ctx, x, w, b, stride, pad, act):
        x = _dense_image(x)
        w = _c(w, "weight")
        N, Cin, H, W = x.shape
        Cout, _, k, _ = w.shape
        col, Ho, Wo = im2col(x, k, stride, pad)
        wp = permute_weight(w)
        y2d = gemm(col, wp, b, N * Ho * Wo, Cout, k * k * Cin, True, True, act)
        ctx.save_for_backward(col, wp, y2d if ACT[act] != C["MMVAE_ACT_NONE"] else None)
        ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, Ho, Wo, act, tuple(w.shape), b is not None,
                   x.is_contiguous() and Cin > 1)
        ctx.leaves = (w, b)
        _tap_relu(w, b, y2d.view(N, Ho, Wo, Cout).permute(0, 3, 1, 2), act)
        return y2d.view(N, Ho, Wo, Cout).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, gy):
        col, wp, y2d = ctx.saved_tensors
        N, Cin, H, W, Cout, k, s, p, Ho, Wo, act, wshape, has_bias, x_nchw = ctx.cfg
        g = act_bwd(_as_rows(_dev(gy, "grad")), y2d, act)
        rows, Kc = N * Ho * Wo, k * k * Cin
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            dcol = gemm(g, wp, None, rows, Kc, Cout, True, False)
            gx = col2im(dcol, None, N, Cin, Ho, Wo, k, s, p, None, nchw_out=x_nchw)
        if ctx.needs_input_grad[1]:
            gwp = gemm(g, col, None, Cout, Kc, rows, False, False)
            dst = _sink(ctx.leaves[0])
            gw = unpermute_weight(gwp, wshape, out=dst)
            gw = None if dst is not None else gw
        if has_bias and ctx.needs_input_grad[2]:
            dst = _sink(ctx.leaves[1])
            gb = colsum(g, out=dst)
            gb = None if dst is not None else gb
        return gx, gw, gb, None, None, None


class ConvT2dAct(torch.autograd.Function):
    """y = act(conv_transpose2d(x, W, b, stride, pad))  (encoders.py:116-136).  W: (Cin,Cout,k,k).
    preact_grad=True: the incoming gradient is already w.r.t. the pre-activation (the consumer fused act').
    mask_input_grad=True: x is the ReLU output of the layer below and the returned gradient is w.r.t. THAT layer's
    pre-activation (its ReLU backward fused here; the layer below must then be built with preact_grad=True).
    Paths: Cout <= 4, k4 s2 p1, NCHW output -> direct kernels (convt_small.cu); otherwise GEMM + col2im lowering."""

    @staticmethod
    def forward(ctx, x, w, b, stride, pad, act, nchw_out, preact_grad, mask_input_grad=False):
        y = ConvT2dAct._forward(ctx, x, w, b, stride, pad, act, nchw_out, preact_grad, mask_input_grad)
        _tap_relu(w, b, y, act)
        return y

    @staticmethod
    def _forward(ctx, x, w, b, stride, pad, act, nchw_out, preact_grad, mask_input_grad):
        x = _dev(x, "x")
        ctx.leaves = (w, b)
        w = _c(w, "weight")
        N, Cin, H, W = x.shape
        _, Cout, k, _ = w.shape
        lib = _lib.load()
        small = (k == 4 and stride == 2 and pad == 1 and nchw_out and w.shape[3] == 4
                 and lib.mmvae_convt_small_supported(H, W, Cin, Cout) == 2)
        x2d = _as_rows(x)
        tc = (ALLOW_TC and not small and k == 4 and stride == 2 and pad == 1 and not nchw_out and w.shape[3] == 4
              and lib.mmvae_convt_tc_supported(H, W, Cin, Cout) == 1)
        if tc and ALLOW_TS and lib.mmvae_convt_ts_supported(H, W, Cin, Cout) == 1:
            def build():
                ne = lib.mmvae_convt_ts_packed_elems(Cin, Cout)
                wf_ = torch.empty(ne, device=x.device, dtype=torch.int16)
                wd_ = torch.empty(ne, device=x.device, dtype=torch.int16)
                _call("mmvae_convt_ts_pack_weights", _p(w), _p(wf_), _p(wd_), Cin, Cout, _stream())
                return wf_, wd_
            wfb, wdb = _layout(w, "ts_bf16", build)
            y = torch.empty((N, 2 * H, 2 * W, Cout), device=x.device, dtype=torch.float32)
            _call("mmvae_convt_ts_fwd", _p(x2d), _p(wfb), _p(b), _p(y), N, H, W, Cin, Cout, ACT[act], _stream(),
                  work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
            keep_y = (ACT[act] != C["MMVAE_ACT_NONE"]) and not preact_grad
            ctx.save_for_backward(x2d, w, y if keep_y else None)
            ctx.wdb = wdb
            ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, act, tuple(w.shape), b is not None, preact_grad,
                       mask_input_grad, "ts")
            return y.permute(0, 3, 1, 2)
        if tc:
            wf = torch.empty((4 * Cout, 4 * Cin), device=x.device, dtype=torch.float32)
            _call("mmvae_convt_weight_classes", _p(w), _p(wf), Cin, Cout, _stream())
            y = torch.empty((N, 2 * H, 2 * W, Cout), device=x.device, dtype=torch.float32)
            _call("mmvae_convt_tc_fwd", _p(x2d), _p(wf), _p(b), _p(y), N, H, W, Cin, Cout, ACT[act], _stream(),
                  work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
            keep_y = (ACT[act] != C["MMVAE_ACT_NONE"]) and not preact_grad
            ctx.save_for_backward(x2d, w, y if keep_y else None)
            ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, act, tuple(w.shape), b is not None, preact_grad,
                       mask_input_grad, "tc")
            return y.permute(0, 3, 1, 2)
        if small:
            y = torch.empty((N, Cout, 2 * H, 2 * W), device=x.device, dtype=torch.float32)
            _call("mmvae_convt_small_fwd", _p(x2d), _p(w), _p(b), _p(y), N, H, W, Cin, Cout, ACT[act], _stream(),
                  work=("byte", 4.0 * (x2d.numel() + y.numel())))
            keep_y = (ACT[act] != C["MMVAE_ACT_NONE"]) and not preact_grad
            ctx.save_for_backward(x2d, w, y if keep_y else None)
            ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, act, tuple(w.shape), b is not None, preact_grad,
                       mask_input_grad, "small")
            return y
        wp = permute_weight(w)                                  # [Cin, k*k*Cout]
        if H == 1 and W == 1 and stride == 1 and pad == 0 and not nchw_out:
            # a 1x1 input makes the layer a linear map onto the k x k x Cout NHWC block (col2im is the identity):
            # one GEMM with the bias (tiled over the k*k taps) and the activation in its epilogue
            bt = b.repeat(k * k) if b is not None else None
            y = gemm(x2d, wp, bt, N, k * k * Cout, Cin, True, False, act=act)
            keep_y = (ACT[act] != C["MMVAE_ACT_NONE"]) and not preact_grad
            ctx.save_for_backward(x2d, wp, y if keep_y else None)
            ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, act, tuple(w.shape), b is not None, preact_grad,
                       mask_input_grad, "linear")
            return y.view(N, k, k, Cout).permute(0, 3, 1, 2)
        col = gemm(x2d, wp, None, N * H * W, k * k * Cout, Cin, True, False)
        y = col2im(col, b, N, Cout, H, W, k, stride, pad, act, nchw_out=nchw_out)
        keep_y = (ACT[act] != C["MMVAE_ACT_NONE"]) and not preact_grad
        ctx.save_for_backward(x2d, wp, y if keep_y else None)
        ctx.cfg = (N, Cin, H, W, Cout, k, stride, pad, act, tuple(w.shape), b is not None, preact_grad,
                   mask_input_grad, "lowered")
        return y

    @staticmethod
    def backward(ctx, gy):
        x2d, wp, y = ctx.saved_tensors
        N, Cin, H, W, Cout, k, s, p, act, wshape, has_bias, preact_grad, mask_in, path = ctx.cfg
        small = path == "small"
        if path in ("tc", "ts"):
            w = wp
            g = _as_rows(_dev(gy, "grad"))                       # [N*2H*2W, Cout], NHWC
            if y is not None:
                g = act_bwd(g, y.view(-1, Cout), act)
            gx = gw = gb = None
            if ctx.needs_input_grad[0] and path == "ts":
                gx2d = torch.empty_like(x2d)
                _call("mmvae_convt_ts_dgrad", _p(g), _p(ctx.wdb), _p(x2d if mask_in else None), _p(gx2d), N, H, W, Cin, Cout,
                      _stream(), work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
                gx = gx2d.view(N, H, W, Cin).permute(0, 3, 1, 2)
            elif ctx.needs_input_grad[0]:
                wpm = permute_weight(w)                          # [Cin, 16*Cout]
                gx2d = torch.empty_like(x2d)
                _call("mmvae_convt_tc_dgrad", _p(g), _p(wpm), _p(x2d if mask_in else None), _p(gx2d), N, H, W, Cin, Cout,
                      _stream(), work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
                gx = gx2d.view(N, H, W, Cin).permute(0, 3, 1, 2)
            need_w, need_b = ctx.needs_input_grad[1], has_bias and ctx.needs_input_grad[2]
            dw, db, sunk = _sinks(ctx.leaves, need_w, need_b)
            lib = _lib.load()
            ts_w = path == "ts" and need_w and Cin <= 128
            if ts_w:
                with _wgrad_side(sunk, g, x2d):
                    nb = lib.mmvae_convt_ts_wgrad_workspace_bytes(N, H, W, Cin, Cout)
                    ws = torch.empty(nb // 4, device=g.device, dtype=torch.float32)
                    gw = torch.empty(wshape, device=g.device, dtype=torch.float32) if dw is None else dw
                    gb = (torch.empty(Cout, device=g.device, dtype=torch.float32) if db is None else db) if need_b else None
                    _call("mmvae_convt_ts_wgrad", _p(x2d), _p(g), _p(gw), _p(gb), _p(ws), nb, N, H, W, Cin, Cout, _stream(),
                          work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
                    gw = None if dw is not None else gw
                    gb = None if db is not None else gb
                return gx, gw, gb, None, None, None, None, None, None
            with _wgrad_side(sunk, g, x2d):
                if need_w:
                    nb = lib.mmvae_convt_tc_wgrad_workspace_bytes(N, H, W, Cin, Cout)
                    ws = torch.empty(nb // 4, device=g.device, dtype=torch.float32) if nb else None
                    gwp = torch.empty((Cin, 16 * Cout), device=g.device, dtype=torch.float32)
                    _call("mmvae_convt_tc_wgrad", _p(x2d), _p(g), _p(gwp), _p(ws), nb, N, H, W, Cin, Cout, _stream(),
                          work=("flop", 2.0 * N * H * W * 16 * Cin * Cout))
                    gw = unpermute_weight(gwp, wshape, out=dw)
                    gw = None if dw is not None else gw
                if need_b:
                    gb = colsum(g, out=db)
                    gb = None if db is not None else gb
            return gx, gw, gb, None, None, None, None, None, None
        if small:
            g = _c(gy, "grad")                                   # (N,Cout,2H,2W) NCHW
            w = wp
            gx = gw = gb = None
            if ctx.needs_input_grad[0]:
                gx2d = torch.empty_like(x2d)
                _call("mmvae_convt_small_dgrad", _p(g), _p(y), _p(w), _p(x2d if mask_in else None), _p(gx2d), N, H, W,
                      Cin, Cout, ACT[act], _stream(), work=("byte", 4.0 * (g.numel() + 2 * x2d.numel())))
                gx = gx2d.view(N, H, W, Cin).permute(0, 3, 1, 2)
            if ctx.needs_input_grad[1] or (has_bias and ctx.needs_input_grad[2]):
                lib = _lib.load()
                need_w, need_b = ctx.needs_input_grad[1], has_bias and ctx.needs_input_grad[2]
                dw, db, sunk = _sinks(ctx.leaves, need_w, need_b)
                with _wgrad_side(sunk and need_w and (need_b or not has_bias), g, x2d, y):
                    ws = torch.empty(lib.mmvae_convt_small_wgrad_workspace_bytes(N, Cin, Cout) // 4, device=g.device,
                                     dtype=torch.float32)
                    gw = torch.empty(wshape, device=g.device, dtype=torch.float32) if dw is None else dw
                    gb = (torch.empty(Cout, device=g.device, dtype=torch.float32) if db is None else db) if has_bias \
                        else None
                    _call("mmvae_convt_small_wgrad", _p(x2d), _p(g), _p(y), _p(gw), _p(gb), _p(ws), N, H, W, Cin, Cout,
                          ACT[act], _stream(), work=("byte", 4.0 * (g.numel() + x2d.numel())))
                gw = None if dw is not None else gw
                gb = None if db is not None else gb
            return gx, gw, gb, None, None, None, None, None, None
        if path == "linear":
            Kc = k * k * Cout
            g = _as_rows(_dev(gy, "grad")).view(N, Kc)            # NHWC rows of the k x k x Cout block
            if y is not None:
                g = act_bwd(g, y.view(N, Kc), act)
            gx = gw = gb = None
            if ctx.needs_input_grad[0]:
                gx2d = gemm(g, wp, None, N, Cin, Kc, True, True)
                if mask_in:
                    gx2d = act_bwd(gx2d, x2d, "relu")
                gx = gx2d.view(N, Cin, 1, 1)
            if ctx.needs_input_grad[1]:
                dst = _sink(ctx.leaves[0])
                gw = unpermute_weight(gemm(x2d, g, None, Cin, Kc, N, False, False), wshape, out=dst)
                gw = None if dst is not None else gw
            if has_bias and ctx.needs_input_grad[2]:
                gb = colsum(g).view(k * k, Cout).sum(0)
            return gx, gw, gb, None, None, None, None, None, None
        gy = _dense_image(gy)
        if not preact_grad and y is not None:
            if gy.stride() != y.stride():
                gy = gy.contiguous(memory_format=torch.channels_last) if y.stride(1) == 1 else gy.contiguous()
            g = act_bwd(gy, y, act)                              # same memory order as y
        else:
            g = gy
        rows, Kc = N * H * W, k * k * Cout
        gx = gw = gb = None
        dcol, _, _ = im2col(g, k, s, p)                          # [N*H*W, k*k*Cout]
        if ctx.needs_input_grad[0]:
            gx2d = gemm(dcol, wp, None, rows, Cin, Kc, True, True)
            if mask_in:
                gx2d = act_bwd(gx2d, x2d, "relu")
            gx = gx2d.view(N, H, W, Cin).permute(0, 3, 1, 2)
        if ctx.needs_input_grad[1]:
            gwp = gemm(x2d, dcol, None, Cin, Kc, rows, False, False)
            dst = _sink(ctx.leaves[0])
            gw = unpermute_weight(gwp, wshape, out=dst)
            gw = None if dst is not None else gw
        if has_bias and ctx.needs_input_grad[2]:
            gb = colsum(_as_rows(g))
        return gx, gw, gb, None, None, None, None, None, None


class RSample(torch.autograd.Function):
    """z[m,k,b,:] = mu[m,b,:] + std[m,b,:] * t(noise[m,k,b,:]);  std from the raw head by `xform`.
    Returns (z [M,K,B,L] (or [M,B,K,L] when bk_layout), std [M,B,L])."""

    @staticmethod
    def forward(ctx, mu, raw, noise, dist, xform, bk_layout):
        mu, raw, noise = _c(mu, "mu"), _c(raw, "raw"), _c(noise, "noise")
        M, B, L = mu.shape
        K = noise.shape[2] if bk_layout else noise.shape[1]
        assert noise.shape == ((M, B, K, L) if bk_layout else (M, K, B, L)), (noise.shape, mu.shape)
        z = torch.empty_like(noise)
        std = torch.empty_like(mu)
        _call("mmvae_rsample_fwd", _p(mu), _p(raw), _p(noise), _p(z), _p(std), M, K, B, L, DIST[dist], XFORM[xform],
              int(bk_layout), _stream())
        ctx.save_for_backward(noise, raw, std)
        ctx.cfg = (M, K, B, L, dist, xform, bk_layout)
        return z, std

    @staticmethod
    def backward(ctx, gz, gstd):
        noise, raw, std = ctx.saved_tensors
        M, K, B, L, dist, xform, bk = ctx.cfg
        gz = _c(gz, "grad_z") if gz is not None else torch.zeros_like(noise)
        gstd = _c(gstd, "grad_std") if gstd is not None else None
        gmu, graw = torch.empty_like(std), torch.empty_like(std)
        _call("mmvae_rsample_bwd", _p(gz), _p(gstd), _p(noise), _p(raw), _p(std), _p(gmu), _p(graw), M, K, B, L,
              DIST[dist], XFORM[xform], int(bk), _stream())
        return gmu, graw, None, None, None, None


class LogLik(torch.autograd.Function):
    """lp[r] = lik_scaling * sum_j log p(x[(r // x_div) % B, j] | recon[r, j]).  recon [R,D], x [B,D].
    Returns (lp, lp_lo): the fp32 value and the residual of the fp64 row sum (non-differentiable; see mmvae_b200.h)."""

    @staticmethod
    def forward(ctx, recon, x, x_div, lik, lik_scaling, through_sigmoid):
        recon, x = _c(recon, "recon"), _c(x, "x")
        R, D = recon.shape
        B = x.shape[0]
        assert x.shape[1] == D
        lp = torch.empty(R, device=recon.device, dtype=torch.float32)
        lo = torch.empty(R, device=recon.device, dtype=torch.float32)
        _call("mmvae_loglik_fwd", _p(recon), _p(x), _p(lp), _p(lo), R, B, D, x_div, LIK[lik], float(lik_scaling), _stream(),
              work=("byte", 4.0 * R * D))
        ctx.save_for_backward(recon, x)
        ctx.cfg = (R, B, D, x_div, lik, float(lik_scaling), through_sigmoid)
        ctx.mark_non_differentiable(lo)
        return lp, lo

    @staticmethod
    def backward(ctx, glp, _glo):
        recon, x = ctx.saved_tensors
        R, B, D, x_div, lik, c, ts = ctx.cfg
        g = torch.empty_like(recon)
        _call("mmvae_loglik_bwd", _p(_c(glp, "grad_lp")), _p(recon), _p(x), _p(g), R, B, D, x_div, LIK[lik], c,
              int(ts), _stream(), work=("byte", 8.0 * R * D))
        return g, None, None, None, None, None


class MoeLogW(torch.autograd.Function):
    """Fused mixture-of-experts objective.  Returns (loss scalar, lw [M,K,B], wt [M,K,B], per-datapoint [B])."""

    @staticmethod
    def forward(ctx, z, mu, std, lpx, lpx_lo, pz_loc, pz_scale, dist, mode):
        z, mu, std, lpx = _c(z, "z"), _c(mu, "mu"), _c(std, "std"), _c(lpx, "lpx")
        lpx_lo = None if lpx_lo is None else _c(lpx_lo, "lpx_lo")
        assert lpx_lo is None or lpx_lo.shape == lpx.shape
        pz_loc, pz_scale = _c(pz_loc.reshape(-1), "pz_loc"), _c(pz_scale.reshape(-1), "pz_scale")
        M, K, B, L = z.shape
        assert lpx.shape == (M, M, K, B), lpx.shape
        lw = torch.empty((M, K, B), device=z.device, dtype=torch.float32)
        wt = torch.empty_like(lw)
        loss = torch.empty((), device=z.device, dtype=torch.float32)
        per = torch.empty(B, device=z.device, dtype=torch.float32)
        _call("mmvae_moe_logw_fwd", _p(z), _p(mu), _p(std), _p(lpx), _p(lpx_lo), _p(pz_loc), _p(pz_scale), _p(lw), _p(wt),
              _p(loss), _p(per), M, K, B, L, DIST[dist], OBJ[mode], _stream())
        ctx.save_for_backward(z, mu, std, wt, pz_loc, pz_scale)
        ctx.cfg = (M, K, B, L, dist, mode)
        ctx.mark_non_differentiable(lw, wt, per)
        return loss, lw, wt, per

    @staticmethod
    def backward(ctx, gloss, _glw, _gwt, _gper):
        z, mu, std, wt, pz_loc, pz_scale = ctx.saved_tensors
        M, K, B, L, dist, mode = ctx.cfg
        g = _c(gloss.reshape(1), "grad_loss")
        gz = torch.empty_like(z)
        gmu, gstd = torch.empty_like(mu), torch.empty_like(std)
        glpx = torch.empty((M, M, K, B), device=z.device, dtype=torch.float32)
        _call("mmvae_moe_logw_bwd", _p(g), _p(z), _p(mu), _p(std), _p(wt), _p(pz_loc), _p(pz_scale), _p(gz), _p(gmu),
              _p(gstd), _p(glpx), M, K, B, L, DIST[dist], OBJ[mode], _stream())
        dreg = mode in ("dreg", "dreg_looser")
        return gz, (None if dreg else gmu), (None if dreg else gstd), glpx, None, None, None, None, None


class Elbo(torch.autograd.Function):
    """m_elbo (objectives.py:87-111).  z is used detached (objectives.py:101).  Returns (loss, terms [M*M], per [B])."""

    @staticmethod
    def forward(ctx, z, mu, std, lpx_diag, pz_loc, pz_scale, dist):
        z, mu, std, lpx_diag = _c(z, "z"), _c(mu, "mu"), _c(std, "std"), _c(lpx_diag, "lpx")
        pz_loc, pz_scale = _c(pz_loc.reshape(-1), "pz_loc"), _c(pz_scale.reshape(-1), "pz_scale")
        M, K, B, L = z.shape
        assert lpx_diag.shape == (M, K, B)
        loss = torch.empty((), device=z.device, dtype=torch.float32)
        per = torch.empty(B, device=z.device, dtype=torch.float32)
        terms = torch.empty(M * M, device=z.device, dtype=torch.float32)
        _call("mmvae_elbo_fwd", _p(z), _p(mu), _p(std), _p(lpx_diag), _p(pz_loc), _p(pz_scale), _p(loss), _p(per),
              _p(terms), M, K, B, L, DIST[dist], _stream())
        ctx.save_for_backward(z, mu, std, lpx_diag, pz_loc, pz_scale)
        ctx.cfg = (M, K, B, L, dist)
        ctx.mark_non_differentiable(terms, per)
        return loss, terms, per

    @staticmethod
    def backward(ctx, gloss, _gt, _gp):
        z, mu, std, lpx_diag, pz_loc, pz_scale = ctx.saved_tensors
        M, K, B, L, dist = ctx.cfg
        g = _c(gloss.reshape(1), "grad_loss")
        gmu, gstd, glpx = torch.empty_like(mu), torch.empty_like(std), torch.empty_like(lpx_diag)
        _call("mmvae_elbo_bwd", _p(g), _p(z), _p(mu), _p(std), _p(lpx_diag), _p(pz_loc), _p(pz_scale), _p(gmu),
              _p(gstd), _p(glpx), M, K, B, L, DIST[dist], _stream())
        return None, gmu, gstd, glpx, None, None, None


# ------------------------------------------------------------------------------------------- flows / JMVAE-NF
class MaskWeight(torch.autograd.Function):
    """mask * W of MaskedLinear (layers.py:24-25); the gradient is masked the same way."""

    @staticmethod
    def forward(ctx, w, mask):
        ctx.save_for_backward(mask)
        return mask_weight(w, mask)

    @staticmethod
    def backward(ctx, g):
        (mask,) = ctx.saved_tensors
        return mask_weight(_c(g, "grad"), mask), None


MADE_FUSED = True         # tests switch it off to compare with the per-layer formulation
MADE_FUSED_MAX_ROWS = 1024   # above this the per-layer GEMMs (tensor cores) win


def _made_arrays(layers):
    """-> (dims int array, [W, mask, b]*n pointer array) for mmvae_made_*: `layers` = [(W, mask, bias or None), ...]."""
    dims = [layers[0][0].shape[1]] + [w.shape[0] for w, _, _ in layers]
    ptrs = []
    for w, m, b in layers:
        ptrs += [w.data_ptr(), m.data_ptr(), b.data_ptr() if b is not None else None]
    return (ctypes.c_int * len(dims))(*dims), (ctypes.c_void_p * len(ptrs))(*ptrs), dims


def made_fused_ok(layers, rows):
    if not MADE_FUSED or rows > MADE_FUSED_MAX_ROWS or any(b is None for _, _, b in layers):
        return False
    dims = [layers[0][0].shape[1]] + [w.shape[0] for w, _, _ in layers]
    return _lib.load().mmvae_made_supported(len(layers), (ctypes.c_int * len(dims))(*dims)) == 1


class MadeFused(torch.autograd.Function):
    """h = MADE(x) with every masked layer in one kernel (csrc/made_train.cu); args: x, then (W, mask, bias) per layer."""

    @staticmethod
    def forward(ctx, x, *params):
        n = len(params) // 3
        x = _c(x, "x")
        layers = [(_c(params[3 * l], "weight"), _c(params[3 * l + 1], "mask"), _c(params[3 * l + 2], "bias")) for l in range(n)]
        dims_c, ptrs_c, dims = _made_arrays(layers)
        R, H = x.shape[0], dims[1]
        h = torch.empty((n - 1, R, H), device=x.device, dtype=torch.float32)
        y = torch.empty((R, dims[-1]), device=x.device, dtype=torch.float32)
        _call("mmvae_made_fwd", _p(x), ptrs_c, dims_c, n, _p(h), _p(y), R, _stream(),
              work=("flop", 2.0 * R * sum(a * b for a, b in zip(dims[:-1], dims[1:]))))
        if RELU_TAP is not None:
            for l in range(n - 1):
                RELU_TAP.setdefault(layers[l][2].data_ptr(), []).append(h[l].detach() > 0)
        ctx.save_for_backward(x, h, *[t for lay in layers for t in lay])
        ctx.leaves = params
        ctx.n = n
        return y

    @staticmethod
    def backward(ctx, gy):
        x, h = ctx.saved_tensors[:2]
        flat = ctx.saved_tensors[2:]
        n = ctx.n
        layers = [(flat[3 * l], flat[3 * l + 1], flat[3 * l + 2]) for l in range(n)]
        dims_c, ptrs_c, dims = _made_arrays(layers)
        R = x.shape[0]
        gy = _c(gy, "grad")
        gpre = torch.empty_like(h)
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        _call("mmvae_made_bwd_data", _p(gy), ptrs_c, dims_c, n, _p(h), _p(gpre), _p(gx), R, _stream(),
              work=("flop", 2.0 * R * sum(a * b for a, b in zip(dims[:-1], dims[1:]))))
        out, dst, sunk_all, any_need = [gx], [], True, False
        for l in range(n):
            need_w, need_b = ctx.needs_input_grad[1 + 3 * l], ctx.needs_input_grad[3 + 3 * l]
            dw = _sink(ctx.leaves[3 * l]) if need_w else None
            db = _sink(ctx.leaves[3 * l + 2]) if need_b else None
            gw = dw if dw is not None else (torch.empty_like(layers[l][0]) if need_w else None)
            gb = db if db is not None else (torch.empty_like(layers[l][2]) if need_b else None)
            sunk_all = sunk_all and (dw is not None or not need_w) and (db is not None or not need_b)
            any_need = any_need or need_w or need_b
            dst += [gw, gb]
            out += [None if dw is not None else gw, None, None if db is not None else gb]
        if any_need:
            grads_c = (ctypes.c_void_p * len(dst))(*[t.data_ptr() if t is not None else None for t in dst])
            with _wgrad_side(sunk_all, gy, gpre, h, x):
                _call("mmvae_made_bwd_weights", _p(x), _p(gy), ptrs_c, grads_c, dims_c, n, _p(h), _p(gpre), R, _stream(),
                      work=("flop", 2.0 * R * sum(a * b for a, b in zip(dims[:-1], dims[1:]))))
        return tuple(out)


class FlowAffine(torch.autograd.Function):
    """Affine coupling + log|det J| update (mmvae_flow_affine_fwd).  -> (out [R,L], ladj_out [R])."""

    @staticmethod
    def forward(ctx, x, h, carry, ladj_in, mode, col, flip_out):
        x, h, ladj_in = _c(x, "x"), _c(h, "h"), _c(ladj_in, "ladj")
        carry = None if carry is None else _c(carry, "carry")
        R, L = x.shape
        assert h.shape == (R, 2 * L) and ladj_in.shape == (R,)
        out, ladj = torch.empty_like(x), torch.empty_like(ladj_in)
        _call("mmvae_flow_affine_fwd", _p(x), _p(h), _p(carry), _p(ladj_in), _p(out), _p(ladj), R, L, int(mode), int(col),
              int(flip_out), _stream())
        ctx.save_for_backward(x, h)
        ctx.cfg = (R, L, int(mode), int(col), int(flip_out), carry is not None)
        return out, ladj

    @staticmethod
    def backward(ctx, g_out, g_ladj):
        x, h = ctx.saved_tensors
        R, L, mode, col, flip_out, has_carry = ctx.cfg
        g_out = _c(g_out, "grad_out") if g_out is not None else torch.zeros_like(x)
        g_ladj = None if g_ladj is None else _c(g_ladj, "grad_ladj")
        g_x, g_h = torch.empty_like(x), torch.empty_like(h)
        g_carry = torch.empty_like(x) if col >= 0 else None
        _call("mmvae_flow_affine_bwd", _p(g_out), _p(g_ladj), _p(x), _p(h), _p(g_x), _p(g_h), _p(g_carry), R, L, mode, col,
              flip_out, _stream())
        return g_x, g_h, (g_carry if has_carry else None), g_ladj, None, None, None


class FlipCols(torch.autograd.Function):
    """x.flip(dims=(1,)) for a 2-D tensor."""

    @staticmethod
    def forward(ctx, x):
        x = _c(x, "x")
        out = torch.empty_like(x)
        _call("mmvae_flip_cols", _p(x), _p(out), x.shape[0], x.shape[1], _stream())
        return out

    @staticmethod
    def backward(ctx, g):
        return FlipCols.apply(g)


GAUSS_KIND = {"std": 0, "log_var": 1, "kl_prior": 2}


class GaussRows(torch.autograd.Function):
    """Row sums of Gaussian log densities / the closed-form prior KL (mmvae_gauss_rows_fwd).  z [R,L] (None for
    'kl_prior'); mu, s [R/div, L].  -> [R].  Differentiable for div == 1."""

    @staticmethod
    def forward(ctx, z, mu, s, kind, div):
        mu, s = _c(mu, "mu"), _c(s, "s")
        z = None if z is None else _c(z, "z")
        L = mu.shape[-1]
        R = mu.shape[0] * div if z is None else z.shape[0]
        assert mu.shape[0] * div == R and s.shape == mu.shape
        out = torch.empty(R, device=mu.device, dtype=torch.float32)
        _call("mmvae_gauss_rows_fwd", _p(z), _p(mu), _p(s), _p(out), R, L, GAUSS_KIND[kind], int(div), _stream())
        ctx.save_for_backward(z, mu, s)
        ctx.cfg = (R, L, kind, div)
        return out

    @staticmethod
    def backward(ctx, g):
        z, mu, s = ctx.saved_tensors
        R, L, kind, div = ctx.cfg
        if div != 1:
            raise RuntimeError("GaussRows is differentiable only for div == 1 (the evaluators run under no_grad)")
        g = _c(g, "grad")
        g_z = None if z is None else torch.empty_like(z)
        g_mu, g_s = torch.empty_like(mu), torch.empty_like(s)
        _call("mmvae_gauss_rows_bwd", _p(g), _p(z), _p(mu), _p(s), _p(g_z), _p(g_mu), _p(g_s), R, L, GAUSS_KIND[kind],
              _stream())
        return g_z, g_mu, g_s, None, None


# ----------------------------------------------------------------------------------- evaluator (no autograd)
QMODE = {"mixture_half": 0, "mixture": 1, "single": 2, "none": 3, "external": 4}


def eval_logw(z, mu, std, lpx, pz_loc, pz_scale, dist, qmode, lq_ext=None):
    """a[b,k] = sum_d lpx[d,b,k] + log p(z[b,k]) - q-term.  z [B,K,L]; mu,std [E,B,L]; lpx [D,B,K]."""
    z, lpx = _c(z, "z"), _c(lpx, "lpx")
    B, K, L = z.shape
    D = lpx.shape[0]
    E = 0 if mu is None else mu.shape[0]
    a = torch.empty((B, K), device=z.device, dtype=torch.float32)
    _call("mmvae_eval_logw", _p(z), _p(None if mu is None else _c(mu, "mu")), _p(None if std is None else _c(std, "std")),
          _p(lpx), _p(None if lq_ext is None else _c(lq_ext, "lq_ext")), _p(_c(pz_loc.reshape(-1), "pz_loc")),
          _p(_c(pz_scale.reshape(-1), "pz_scale")), _p(a), E, D, B, K, L, DIST[dist], QMODE[qmode], _stream())
    return a


def lse_partial(a, mx=None, se=None):
    """merge the K-shard a [B,Kc] into the running per-datapoint (max, sum exp) pair."""
    a = _c(a, "a")
    B, Kc = a.shape
    acc = mx is not None
    if not acc:
        mx = torch.empty(B, device=a.device, dtype=torch.float32)
        se = torch.empty(B, device=a.device, dtype=torch.float32)
    _call("mmvae_lse_partial", _p(a), _p(mx), _p(se), B, Kc, int(acc), _stream())
    return mx, se


def lse_finish(mx, se, K_total):
    ll = torch.empty_like(mx)
    _call("mmvae_lse_finish", _p(mx), _p(se), _p(ll), mx.numel(), float(math.log(K_total)), _stream())
    return ll


def mask_weight(w, mask):
    out = torch.empty_like(w)
    _call("mmvae_mask_weight", _p(_c(w, "w")), _p(_c(mask, "mask")), _p(out), w.numel(), _stream())
    return out


def adam_amsgrad_(p, g, m, v, vmax, step_dev, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, amsgrad=True):
    """in-place fused Adam(amsgrad) on flat buffers (main.py:114-115)."""
    weights_changed()
    _call("mmvae_adam_amsgrad", _p(p), _p(g), _p(m), _p(v), _p(vmax), p.numel(), float(lr), float(betas[0]),
          float(betas[1]), float(eps), _p(step_dev), int(amsgrad), _stream())


# ------------------------------------------------------------------------- fused sequential MADE direction (inference)
class MadeSeqPlan:
    """Degree-sorted packing of one MADE (flows.MADE with three equal hidden layers and the reference's sequential degrees,
    made_model.py:68-95) for mmvae_made_seq_fwd: the gather index that turns (biases | masked weights) into the pack of
    csrc/made_seq.cu, and the per-degree unit counts.  Built once per MADE; `ok` is False when the kernel does not apply
    (other depth, more than 8 units per degree, masks that are not the sequential ones, shared memory) — the caller then
    keeps the GEMM formulation."""

    def __init__(self, made):
        import numpy as np
        lay = made.layers()
        L, hs = made.input_dim, made.hidden_sizes
        self.ok = False
        if len(hs) != 3 or len(set(hs)) != 1 or made.output_dim != 2 * L or not (2 <= L <= 32):
            return
        H = hs[0]
        from .flows import made_masks
        std = made_masks(L, hs, 2 * L)
        if any(tuple(l.mask.shape) != tuple(m.shape) or not torch.equal(l.mask.detach().cpu(), m) for l, m in zip(lay, std)):
            return
        deg = np.arange(H) % (L - 1)
        perm = np.argsort(deg, kind="stable")
        G = L - 1
        cnt = np.array([(deg <= d).sum() for d in range(G)], dtype=np.int32)
        c0 = np.concatenate([[0], cnt])                       # c0[d] = units of degree < d
        if (np.diff(c0) > 8).any():
            return
        lib = _lib.load()
        self.cnt_host = (ctypes.c_int * G)(*[int(v) for v in cnt])
        n = lib.mmvae_made_seq_pack_floats(ctypes.cast(self.cnt_host, ctypes.c_void_p), L)
        if n <= 0 or lib.mmvae_made_seq_smem_bytes(ctypes.cast(self.cnt_host, ctypes.c_void_p), L, H) > 227 * 1024:
            return
        ob = [0, H, 2 * H, 3 * H]
        oW1 = 3 * H + 2 * L
        oW2, oW3 = oW1 + H * L, oW1 + H * L + H * H
        oW4 = oW3 + H * H
        zero = oW4 + 2 * L * H
        idx = np.full(n, zero, dtype=np.int64)
        pos = 0
        for layer in range(3):                                # b1g, b2g, b3g [G][8]
            for d in range(G):
                for u in range(c0[d + 1] - c0[d]):
                    idx[pos + 8 * d + u] = ob[layer] + perm[c0[d] + u]
            pos += 8 * G
        idx[pos:pos + 2 * L] = ob[3] + np.arange(2 * L)       # b4
        pos += (2 * L + 3) & ~3
        for d in range(G):                                    # W1g: [d+1][8]
            for u in range(c0[d + 1] - c0[d]):
                j = perm[c0[d] + u]
                idx[pos + 8 * np.arange(d + 1) + u] = oW1 + j * L + np.arange(d + 1)
            pos += 8 * (d + 1)
        for oW in (oW2, oW3):                                 # W2g, W3g: [cnt[d]][8]
            for d in range(G):
                kk = perm[:c0[d + 1]]
                for u in range(c0[d + 1] - c0[d]):
                    j = perm[c0[d] + u]
                    idx[pos + 8 * np.arange(c0[d + 1]) + u] = oW + j * H + kk
                pos += 8 * c0[d + 1]
        for i in range(L):                                    # W4g: [cnt[i-1]][2]
            kk = perm[:c0[i]]
            for c in range(2):
                idx[pos + 2 * np.arange(c0[i]) + c] = oW4 + (c * L + i) * H + kk
            pos += 2 * c0[i]
        assert pos <= n
        self.L, self.H, self.n = L, H, int(n)
        self.idx_cpu = torch.from_numpy(idx)
        self.cnt_cpu = torch.from_numpy(cnt.copy())
        self._dev = {}
        self.ok = True

    def on(self, device):
        if device not in self._dev:
            self._dev[device] = (self.idx_cpu.to(device), self.cnt_cpu.to(device))
        return self._dev[device]


def made_seq_plan(made):
    plan = getattr(made, "_seq_plan", None)
    if plan is None:
        plan = MadeSeqPlan(made)
        object.__setattr__(made, "_seq_plan", plan)
    return plan


def made_seq(y, ladj, made, mw, mode, flip_in, flip_out):
    """One flow block in the sequential direction, fused (inference only).  y [R,L], ladj [R]; mw = made.masked_weights()."""
    plan = made_seq_plan(made)
    y, ladj = _c(y, "y"), _c(ladj, "ladj")
    R, L = y.shape
    idx, cnt_dev = plan.on(y.device)
    lay = made.layers()
    src = torch.cat([l.bias.detach().reshape(-1) for l in lay] + [w.detach().reshape(-1) for w in mw] +
                    [torch.zeros(1, device=y.device)])
    pack = src.index_select(0, idx)
    out, ladj_out = torch.empty_like(y), torch.empty_like(ladj)
    _call("mmvae_made_seq_fwd", _p(y), _p(ladj), _p(out), _p(ladj_out), _p(pack), ctypes.c_void_p(cnt_dev.data_ptr()),
          ctypes.cast(plan.cnt_host, ctypes.c_void_p), R, L, plan.H, int(mode), int(flip_in), int(flip_out), _stream())
    return out, ladj_out
