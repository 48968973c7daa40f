"""mmvae_colsum_static: the allocation-free variant of mmvae_colsum (scratch in a file-scope __device__ array).

Collected last on purpose (file name): a failure here must not hide other tests.
Calls go straight through the C ABI (ctypes), outputs are compared with the existing entry point's.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _both(lib, A):
    from mmvae_b200 import _lib
    M, N = A.shape
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = torch.empty(lib.mmvae_colsum_workspace_bytes(M, N) // 4, device=A.device, dtype=torch.float32)
    ref = torch.empty(N, device=A.device, dtype=torch.float32)
    _lib.check(lib.mmvae_colsum(_ptr(A), _ptr(ref), _ptr(ws), M, N, st), "mmvae_colsum")
    out = torch.full((N,), float("nan"), device=A.device, dtype=torch.float32)
    _lib.check(lib.mmvae_colsum_static(_ptr(A), _ptr(out), M, N, st), "mmvae_colsum_static")
    return out, ref


def test_colsum_static_matches_colsum_twice_in_a_row():
    from mmvae_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cpu").manual_seed(11)
    # two calls in a row with different shapes: the second reads a scratch array that still holds the first call's partials
    for (M, N) in ((7680, 512), (1000, 77), (128, 3072)):
        A = torch.randn(M, N, generator=g).cuda()
        out, ref = _both(lib, A)
        torch.cuda.synchronize()
        assert torch.equal(out, ref), (M, N, (out - ref).abs().max().item())      # same kernels, same order: bit-identical
        exact = A.double().sum(0)
        assert ((out.double() - exact).norm() / exact.norm()).item() < 1e-6       # the tolerance test_colsum_and_act_bwd uses
    # same input twice: identical both times
    A = torch.randn(4096, 200, generator=g).cuda()
    o1, _ = _both(lib, A)
    o2, ref = _both(lib, A)
    assert torch.equal(o1, o2) and torch.equal(o1, ref)


def test_colsum_static_rejects_sizes_beyond_the_scratch():
    from mmvae_b200 import _lib
    lib = _lib.load()
    cap = lib.mmvae_colsum_static_max_bytes()
    assert cap == 1 << 20
    M, N = 1024, 300000                                 # one slab of 300000 partial sums > 262144 floats
    assert lib.mmvae_colsum_workspace_bytes(M, N) > cap
    A = torch.empty(1, device="cuda")                    # never dereferenced: the size check comes first
    out = torch.empty(N, device="cuda")
    rc = lib.mmvae_colsum_static(_ptr(A), _ptr(out), M, N, None)
    assert rc == 1
    msg = lib.mmvae_last_error().decode()
    assert "scratch" in msg and str(cap) in msg, msg
