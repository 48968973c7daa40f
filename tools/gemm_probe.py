"""GPU probes of the GEMM entry point (mmvae_gemm: tcgen05 / skinny / SIMT dispatch), one file, one sub-command each:

    python tools/gemm_probe.py shapes [--eval]   the GEMM shapes of a C2 / C3 step: default dispatch vs exact-fp32 SIMT, and the
                                                 latent-sized (K or N = 20) shapes under the three tcgen05 gates
    python tools/gemm_probe.py ksweep            time against K for a few (M, N): looks for a fixed cost
    python tools/gemm_probe.py fixed             ablation of the fixed cost of one tcgen05 launch on small shapes
    python tools/gemm_probe.py ablate [bits..]   ablation of gemm_tc_kernel in its convolution dgrad / wgrad modes
    python tools/gemm_probe.py bf16              bf16x3 (debug bit 65536) against 3xTF32: time and error
    python tools/gemm_probe.py one M N K         one shape launched six times (for ncu --set full --import-source on)

Ablation bits (mmvae_gemm_tc_set_mode(mode | bits << 8); results wrong by construction): 1 no epilogue stores, 2|4 no operand
loads, 8 no hi/lo split, 16 one MMA only, 64 no global store instruction; 32768 = round-1 dispatch gate, 128 / 256 = strict /
K<32 gates, 8192 = 3 UMMAs per k-step, 65536 = bf16x3."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                   # noqa: E402
import mmvae_b200                              # noqa: E402,F401
from mmvae_b200 import _lib, ops               # noqa: E402

lib = _lib.load()
p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def bench(fn, n=20, graph=False):
    """GPU microseconds per call; graph=True replays n captured calls (no host launch cost)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    inner = 1
    if graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n):
                fn()
        fn, inner, n = g.replay, n, 5
        fn()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * inner) * 1e3


def operands(M, N, K, ak=1, bk=1):
    return (torch.randn((M, K) if ak else (K, M), device="cuda"), torch.randn((N, K) if bk else (K, N), device="cuda"),
            torch.randn(N, device="cuda"), torch.empty(M, N, device="cuda"))


def cmd_shapes(argv):
    step = [(128, 512, 784, 1, 1, "enc L1 fwd"), (128, 784, 512, 1, 0, "enc L1 dgrad"), (512, 784, 128, 0, 0, "enc L1 wgrad"),
            (128, 512, 512, 1, 1, "hidden fwd"), (512, 512, 128, 0, 0, "hidden wgrad"), (128, 20, 512, 1, 1, "head fwd"),
            (128, 512, 20, 1, 0, "head dgrad"), (20, 512, 128, 0, 0, "head wgrad"), (128, 128, 20, 1, 1, "MADE in fwd"),
            (128, 128, 128, 1, 1, "MADE hid fwd"), (128, 128, 128, 1, 0, "MADE hid dgrad"), (128, 128, 128, 0, 0, "MADE hid wgrad"),
            (128, 40, 128, 1, 1, "MADE out fwd"), (40, 128, 128, 0, 0, "MADE out wgrad"), (128, 2048, 20, 1, 0, "SVHN dec L1 fwd"),
            (128, 784, 512, 1, 1, "MLP dec L2 fwd"), (2048, 20, 128, 0, 0, "SVHN enc head wgrad"), (128, 40, 2048, 1, 1, "SVHN enc heads fwd"),
            (2048, 2048, 128, 0, 0, "conv wgrad-like"), (8192, 64, 512, 1, 1, "conv lowered L2"), (32768, 32, 48, 1, 1, "enc conv1 lowered"),
            (2048, 128, 1024, 1, 1, "enc conv3 lowered"), (1024, 512, 512, 1, 1, "M=1024"), (4096, 512, 512, 1, 1, "M=4096"),
            (512, 512, 2048, 0, 0, "wgrad K=2048"), (512, 512, 8192, 0, 0, "wgrad K=8192"), (64, 32, 32768, 0, 0, "conv1 wgrad")]
    for M, N, K, ak, bk, what in step:
        A, B, bias, out = operands(M, N, K, ak, bk)
        ts = []
        for tc in (True, False):
            ops.ALLOW_TC = tc
            ts.append(bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out), n=50, graph=True))
        ops.ALLOW_TC = True
        lim = lib.mmvae_gemm_small_set_limit(1 << 40)                              # the cluster split-K kernel, whatever the size
        t_small = bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out), n=50, graph=True)
        lib.mmvae_gemm_small_set_max_split(8)
        t_small8 = bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out), n=50, graph=True)
        lib.mmvae_gemm_small_set_max_split(4)
        t_small4 = bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out), n=50, graph=True)
        lib.mmvae_gemm_small_set_max_split(16)
        lib.mmvae_gemm_small_set_limit(0)
        ts = [bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out), n=50, graph=True), ts[0]]
        lib.mmvae_gemm_small_set_limit(lim)
        At, Bt = (A if ak else A.t()), (B.t() if bk else B)                     # yardstick only: cuBLAS through torch
        ts.append(bench(lambda: torch.relu_(torch.addmm(bias, At, Bt, out=out)), n=50, graph=True))
        print(f"{what:22s} M={M:5d} N={N:5d} K={K:5d} a_k={ak} b_k={bk}: cluster split-K {t_small:6.1f} us (S<=8 {t_small8:5.1f}, S<=4 {t_small4:5.1f})   without it {ts[0]:6.1f} us   "
              f"default dispatch {ts[1]:6.1f} us   (torch addmm+relu {ts[2]:6.1f} us)", flush=True)
    latent = [(7680, 512, 20, 1, 1, "MNIST dec L0 fwd"), (7680, 20, 512, 1, 0, "MNIST dec L0 dgrad"), (512, 20, 7680, 0, 0, "MNIST dec L0 wgrad"),
              (7680, 2048, 20, 1, 0, "SVHN dec L1 fwd"), (7680, 20, 2048, 1, 1, "SVHN dec L1 dgrad"), (20, 2048, 7680, 0, 0, "SVHN dec L1 wgrad"),
              (128, 20, 512, 1, 1, "enc head fwd"), (20, 512, 128, 0, 0, "enc head wgrad"),
              (100000, 512, 20, 1, 1, "eval MNIST dec L0"), (32000, 2048, 20, 1, 0, "eval SVHN dec L1")]
    gates = ((128, "strict gate"), (0, "thin M|N on tcgen05"), (256, "+ K<32 on tcgen05"))
    for M, N, K, ak, bk, what in latent:
        A, B, _, out = operands(M, N, K, ak, bk)
        row = []
        for dbg, tag in gates:
            lib.mmvae_gemm_tc_set_mode(dbg << 8)
            row.append(f"{tag} {bench(lambda: ops.gemm(A, B, None, M, N, K, ak, bk, out=out)):7.1f} us")
        print(f"{what:22s} M={M:6d} N={N:5d} K={K:5d} a_k={ak} b_k={bk}:  " + "   ".join(row), flush=True)
    lib.mmvae_gemm_tc_set_mode(0)
    if "--eval" in argv:
        sys.argv = sys.argv[:1]
        import bench as Bn
        dev = torch.device("cuda", 0)
        model = Bn.make_model(dev)
        for dbg, tag in gates:
            lib.mmvae_gemm_tc_set_mode(dbg << 8)
            ev = Bn.run_eval(model, dev, 64, Bn.peaks())
            print(f"eval leg, {tag}: {ev['value']:.0f} datapoints/s ({ev['s_per_pass'] * 1e3:.2f} ms/pass)", flush=True)
        lib.mmvae_gemm_tc_set_mode(0)


def cmd_ksweep(argv):
    lib.mmvae_gemm_tc_set_mode(32768 << 8)
    for M, N, ak, bk in ((128, 512, 1, 1), (512, 512, 0, 0), (2048, 2048, 1, 1), (7680, 512, 1, 1)):
        row = []
        for K in (32, 64, 96, 128, 160, 192, 256, 512, 1024, 2048):
            A, B, bias, out = operands(M, N, K, ak, bk)
            row.append(f"K={K}: {bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act='relu', out=out), graph=True):.1f}")
        print(f"M={M} N={N} a_k={ak} b_k={bk}  " + "  ".join(row), flush=True)
    lib.mmvae_gemm_tc_set_mode(0)


def cmd_fixed(argv):
    for M, N, K in ((128, 512, 64), (128, 128, 64), (128, 32, 64), (2048, 2048, 64)):
        A, B, bias, out = operands(M, N, K)
        row = []
        for tag, mode, b, act in (("full", 0, bias, "relu"), ("no bias/act", 0, None, None), ("no stores", 1 << 8, bias, "relu"),
                                  ("no global store instr", 64 << 8, bias, "relu"), ("no split", 8 << 8, bias, "relu"),
                                  ("one MMA", 16 << 8, bias, "relu"), ("1xTF32", 1, bias, "relu"),
                                  ("no stores+no split+one MMA", (1 | 8 | 16) << 8, bias, "relu")):
            lib.mmvae_gemm_tc_set_mode(mode | (32768 << 8))
            row.append(f"{tag}: {bench(lambda: ops.gemm(A, B, b, M, N, K, 1, 1, act=act, out=out), graph=True):.1f}")
        lib.mmvae_gemm_tc_set_mode(0)
        print(f"M={M} N={N} K={K}  " + " | ".join(row), flush=True)


def cmd_ablate(argv):
    N = 7680
    names = {1: "noepi", 6: "noload", 8: "noconv", 16: "nomma", 8192: "3 UMMAs per k-step (no hi*[hi|lo] merge)"}
    for (H, Cin, Cout) in ((4, 128, 64), (8, 64, 32)):
        x = torch.randn(N, H, H, Cin, device="cuda")
        w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.1
        g = torch.randn(N, 2 * H, 2 * H, Cout, device="cuda")
        wp, gx, gwp = ops.permute_weight(w), torch.empty_like(x), torch.empty(Cin, 16 * Cout, device="cuda")
        nb = lib.mmvae_convt_tc_wgrad_workspace_bytes(N, H, H, Cin, Cout)
        ws = torch.empty(max(nb // 4, 1), device="cuda")
        for mode in (0, 1):
            for dbg in ([int(a) for a in argv] or (0, 8192, 1, 6, 8, 16, 1 | 8, 1 | 6 | 8, 1 | 6 | 8 | 16)):
                lib.mmvae_gemm_tc_set_mode(mode | (dbg << 8))
                td = bench(lambda: _lib.check(lib.mmvae_convt_tc_dgrad(p(g), p(wp), p(x), p(gx), N, H, H, Cin, Cout, st())), n=10)
                tw = bench(lambda: _lib.check(lib.mmvae_convt_tc_wgrad(p(x), p(g), p(gwp), p(ws), nb, N, H, H, Cin, Cout, st())), n=10)
                tag = " ".join(v for k, v in names.items() if dbg & k)
                print(f"H={H} Cin={Cin} Cout={Cout} passes={3 if mode == 0 else 1} dbg={dbg:4d} [{tag}]: dgrad {td:7.1f} us  wgrad {tw:7.1f} us", flush=True)
    lib.mmvae_gemm_tc_set_mode(0)


def cmd_bf16(argv):
    for (M, N, K) in ((7680, 784, 512), (512, 20, 7680), (8192, 64, 512), (2048, 128, 1024), (7680, 512, 512), (122880, 128, 512)):
        A = torch.randn(M, K, device="cuda") * K ** -0.5
        B, b = torch.randn(N, K, device="cuda"), torch.randn(N, device="cuda")
        ref = A.double() @ B.double().t() + b.double()
        for bits, name in ((65536, "bf16x3"), (0, "3xTF32")):
            lib.mmvae_gemm_tc_set_mode(bits << 8)
            n0 = lib.mmvae_gemm_tc_launch_count()
            out = ops.gemm(A, B, b, M, N, K, 1, 1, None)
            on_tc = lib.mmvae_gemm_tc_launch_count() != n0
            err = ((out.double() - ref).norm() / ref.norm()).item()
            print(f"M={M} N={N} K={K} {name}: {bench(lambda: ops.gemm(A, B, b, M, N, K, 1, 1, None)):7.1f} us  err {err:.2e}  tcgen05={on_tc}", flush=True)
        lib.mmvae_gemm_tc_set_mode(0)


def cmd_one(argv):
    M, N, K = (int(v) for v in argv[:3])
    lib.mmvae_gemm_tc_set_mode(32768 << 8)
    A, B, bias, out = operands(M, N, K)
    for _ in range(6):
        ops.gemm(A, B, bias, M, N, K, 1, 1, act="relu", out=out)
    torch.cuda.synchronize()


if __name__ == "__main__":
    cmds = {"shapes": cmd_shapes, "ksweep": cmd_ksweep, "fixed": cmd_fixed, "ablate": cmd_ablate, "bf16": cmd_bf16, "one": cmd_one}
    if len(sys.argv) < 2 or sys.argv[1] not in cmds:
        sys.exit(__doc__)
    cmds[sys.argv[1]](sys.argv[2:])
