"""Times the last decoder layer's forward (Cin=32 -> Cout=3, 16x16 -> 32x32, 7680 images): tcgen05 lowered-GEMM kernel
(convt_l4_tc.cu) against the direct SIMT kernel (debug bit 512 forces it)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200  # noqa: F401
from mmvae_b200 import _lib
lib = _lib.load()
p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def bench(fn, n=20):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for N in (7680, 32768):
    x = torch.randn(N, 16, 16, 32, device="cuda").relu_()
    w = torch.randn(32, 3, 4, 4, device="cuda") * 0.1
    b = torch.randn(3, device="cuda")
    y = torch.empty(N, 3, 32, 32, device="cuda")
    mb = (x.numel() + y.numel()) * 4 / 1e6
    for dbg, tag in ((512, "SIMT direct"), (0, "tcgen05 lowered"), (1024, "tc, no scatter"), (2048, "tc, no finalize"),
                     (1024 | 2048, "tc, no scatter/finalize"), (4096, "tc, no hi/lo split"), (7168, "tc, MMA + TMA only")):
        lib.mmvae_gemm_tc_set_mode(dbg << 8)
        t = bench(lambda: _lib.check(lib.mmvae_convt_small_fwd(p(x), p(w), p(b), p(y), N, 16, 16, 32, 3, 2, st())))
        print(f"N={N:6d} {tag:28s}: {t:7.1f} us   {mb / t:6.2f} TB/s algorithmic ({mb:.0f} MB)", flush=True)
    g = torch.randn(N, 3, 32, 32, device="cuda")
    gx = torch.empty_like(x)
    mb = (2 * y.numel() + 2 * x.numel()) * 4 / 1e6
    for dbg, tag in ((512, "dgrad SIMT direct"), (0, "dgrad tcgen05"), (1024, "dgrad tc, no epilogue ld/st"), (2048, "dgrad tc, no gather"),
                     (4096, "dgrad tc, no g_pre image"), (2048 | 4096, "dgrad tc, no build"), (1024 | 2048, "dgrad tc, no epi, no gather"),
                     (1024 | 4096, "dgrad tc, no epi, no g_pre"), (7168, "dgrad tc, skeleton")):
        lib.mmvae_gemm_tc_set_mode(dbg << 8)
        t = bench(lambda: _lib.check(lib.mmvae_convt_small_dgrad(p(g), p(y), p(w), p(x), p(gx), N, 16, 16, 32, 3, 2, st())))
        print(f"N={N:6d} {tag:28s}: {t:7.1f} us   {mb / t:6.2f} TB/s algorithmic ({mb:.0f} MB)", flush=True)
    gw = torch.empty_like(w); gb = torch.empty(3, device="cuda")
    ws = torch.empty(lib.mmvae_convt_small_wgrad_workspace_bytes(N, 32, 3) // 4, device="cuda")
    mb = (2 * y.numel() + x.numel()) * 4 / 1e6
    for dbg, tag in ((32, "wgrad SIMT direct"), (0, "wgrad tensor-memory (bf16x3)"), (2, "wgrad ts, no g reads"), (4, "wgrad ts, no x split"),
                     (8, "wgrad ts, no MMAs"), (14, "wgrad ts, skeleton"), (14 | 16, "wgrad ts, skeleton, no g/y loads"), (14 | 64, "wgrad ts, skeleton, no build"),
                     (14 | 16 | 64, "wgrad ts, skeleton, neither"), (16384, "wgrad tcgen05 (1st gen)"), (16384 | 1024, "wgrad tc, no gather"),
                     (16384 | 2048, "wgrad tc, no g_pre image"), (16384 | 4096, "wgrad tc, 1/4 of the UMMAs"),
                     (16384 | 7168, "wgrad tc, skeleton")):
        lib.mmvae_gemm_tc_set_mode(dbg << 8)
        t = bench(lambda: _lib.check(lib.mmvae_convt_small_wgrad(p(x), p(g), p(y), p(gw), p(gb), p(ws), N, 16, 16, 32, 3, 2, st())))
        print(f"N={N:6d} {tag:28s}: {t:7.1f} us   {mb / t:6.2f} TB/s algorithmic ({mb:.0f} MB)", flush=True)
lib.mmvae_gemm_tc_set_mode(0)
