"""GPU probe of the tcgen05 GEMM: layouts, tails, split-K, and whether the tensor core truncates or rounds fp32->tf32."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mmvae_b200
from mmvae_b200 import ops, _lib

lib = _lib.load()
print("has_tcgen05", lib.mmvae_has_tcgen05(), flush=True)
torch.manual_seed(0)


def run(M, N, K, a_k, b_k, mode, bias=False, act=None):
    A = torch.randn(M, K, device="cuda")
    B = torch.randn(N, K, device="cuda")
    bs = torch.randn(N, device="cuda") if bias else None
    a_s = A if a_k else A.t().contiguous()
    b_s = B if b_k else B.t().contiguous()
    lib.mmvae_gemm_tc_set_mode(mode)
    ops.ALLOW_TC = True
    out = ops.gemm(a_s, b_s, bs, M, N, K, a_k, b_k, act)
    torch.cuda.synchronize()
    lib.mmvae_gemm_tc_set_mode(0)
    ref = A.double() @ B.double().t()
    if bias:
        ref = ref + bs.double()
    if act == "relu":
        ref = ref.relu()
    e = ((out.double() - ref).norm() / ref.norm()).item()
    if mode == 1:
        tr = lambda t: (t.view(torch.int32) & -8192).view(torch.float32)
        rt = tr(A).double() @ tr(B).double().t()
        e_tr = ((out.double() - rt).norm() / rt.norm()).item()
        return e, e_tr
    return e, None


cases = [(256, 128, 64), (128, 128, 32), (384, 256, 96), (300, 200, 100), (7680, 784, 512), (1024, 128, 4096),
         (64, 512, 8192), (32, 48, 16384), (122880, 1024, 128), (128, 20, 512)]
for mode in (1, 0):
    for (M, N, K) in cases:
        for a_k in (1, 0):
            for b_k in (1, 0):
                if (not a_k and M % 4) or (not b_k and N % 4):
                    continue
                t0 = time.time()
                e, e_tr = run(M, N, K, a_k, b_k, mode, bias=(M % 3 == 0), act="relu" if N == 128 else None)
                print(f"mode={mode} M={M} N={N} K={K} a_k={a_k} b_k={b_k}: rel_err={e:.3e}"
                      + (f" vs_trunc={e_tr:.3e}" if e_tr is not None else "") + f"  ({time.time()-t0:.2f}s)", flush=True)
print("done")
