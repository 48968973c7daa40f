"""Times the GEMM shapes of a 128-row step (encoders at C2, everything at C3) on the tcgen05 path and on the exact-fp32 SIMT
kernel: back-to-back launches on one stream, i.e. max(launch overhead, kernel duration) per call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import ops


def bench(fn, n=50):
    """GPU time per call: n calls captured in one CUDA graph (no host launch cost)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    fn = g.replay
    n_in = n
    n = 5
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * n_in) * 1e3


if __name__ == "__main__":
    SHAPES = [(128, 512, 784, 1, 1, "enc L1 fwd"), (128, 784, 512, 1, 0, "enc L1 dgrad"), (512, 784, 128, 0, 0, "enc L1 wgrad"),
              (128, 512, 512, 1, 1, "hidden fwd"), (512, 512, 128, 0, 0, "hidden wgrad"),
              (128, 20, 512, 1, 1, "head fwd"), (128, 512, 20, 1, 0, "head dgrad"), (20, 512, 128, 0, 0, "head wgrad"),
              (128, 128, 20, 1, 1, "MADE in fwd"), (128, 128, 128, 1, 1, "MADE hid fwd"), (128, 128, 128, 1, 0, "MADE hid dgrad"),
              (128, 128, 128, 0, 0, "MADE hid wgrad"), (128, 40, 128, 1, 1, "MADE out fwd"), (40, 128, 128, 0, 0, "MADE out wgrad"),
              (128, 2048, 20, 1, 0, "SVHN dec L1 fwd"), (128, 784, 512, 1, 1, "MLP dec L2 fwd"), (2048, 20, 128, 0, 0, "SVHN enc head wgrad"),
              (128, 40, 2048, 1, 1, "SVHN enc heads fwd"), (2048, 2048, 128, 0, 0, "conv wgrad-like"), (8192, 64, 512, 1, 1, "conv lowered L2"),
              (32768, 32, 48, 1, 1, "enc conv1 lowered"), (2048, 128, 1024, 1, 1, "enc conv3 lowered"), (512, 128, 1024, 1, 1, "x"),
              (1024, 512, 512, 1, 1, "M=1024"), (2048, 512, 512, 1, 1, "M=2048"), (4096, 512, 512, 1, 1, "M=4096"),
              (512, 512, 2048, 0, 0, "wgrad K=2048"), (512, 512, 8192, 0, 0, "wgrad K=8192"), (64, 32, 32768, 0, 0, "conv1 wgrad")]
    for M, N, K, ak, bk, what in SHAPES:
        A = torch.randn((M, K) if ak else (K, M), device="cuda")
        B = torch.randn((N, K) if bk else (K, N), device="cuda")
        bias = torch.randn(N, device="cuda")
        out = torch.empty(M, N, device="cuda")
        ts = []
        for tc in (True, False):
            ops.ALLOW_TC = tc
            ts.append(bench(lambda: ops.gemm(A, B, bias, M, N, K, ak, bk, act="relu", out=out)))
        ops.ALLOW_TC = True
        print(f"{what:22s} M={M:5d} N={N:5d} K={K:5d} a_k={ak} b_k={bk}: default path {ts[0]:6.1f} us   SIMT {ts[1]:6.1f} us", flush=True)
