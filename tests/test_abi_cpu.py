"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/mmvae_b200.h declares; the
host-side API rejects anything that would need a fallback.  No compute calls (there is no GPU here)."""
import ctypes
import os

import pytest
import torch


def test_library_exports_every_declared_symbol():
    import __graft_entry__
    __graft_entry__.build()
    from mmvae_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    protos, consts = _lib.parse_header()
    assert len(protos) >= 24
    for name in protos:
        assert hasattr(lib, name), name
    assert _lib.load().mmvae_abi_version() == consts["MMVAE_ABI_VERSION"]
    assert _lib.load().mmvae_last_error() is not None


def test_argument_errors_are_reported_not_swallowed():
    from mmvae_b200 import _lib
    lib = _lib.load()
    rc = lib.mmvae_gemm(None, None, None, None, 4, 4, 4, 1, 1, 0, None, 0, 0, None)     # null pointers: no launch
    assert rc == 1 and b"null" in lib.mmvae_last_error()
    with pytest.raises(RuntimeError, match="gemm"):
        _lib.check(rc, "mmvae_gemm")


def test_no_cpu_fallback_on_the_product_path():
    import mmvae_b200
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mmvae_b200.ops.LinearAct.apply(torch.zeros(2, 3), torch.zeros(4, 3), None, None)
    with pytest.raises(TypeError):
        mmvae_b200.objectives.m_dreg_looser(torch.nn.Linear(2, 2), [torch.zeros(1, 2)], 2)


def test_product_never_imports_the_oracle():
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mmvae_b200")
    for dp, _, fs in os.walk(root):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src.replace("# oracle", ""), f"{f} mentions the oracle"


def test_model_keys_match_reference_state_dict():
    import argparse
    import mmvae_b200
    from oracle import bivae_oracle as O
    from oracle import synth
    p = argparse.Namespace(latent_dim=20, dist="laplace", recon_losses=["normal", "normal"], llik_scaling=0.0)
    m = mmvae_b200.VAE_mnist_svhn(p)
    sd = synth.make_state_dict(O.spec_mnist_svhn(), 1)
    assert set(m.state_dict()) == set(sd)
    assert sum(q.numel() for q in m.parameters()) == 1289547
    assert m.lik_scaling == (3072 / 784, 1.0)
    p3 = argparse.Namespace(latent_dim=30, dist="normal", recon_losses=["normal"] * 3, llik_scaling=0.0)
    m3 = mmvae_b200.VAE_mmvae_msf(p3)
    assert set(m3.state_dict()) == set(synth.make_state_dict(O.spec_msf(), 1))


def test_microbatch_split_matches_reference_rule():
    from mmvae_b200.objectives import compute_microbatch_split
    x = [torch.zeros(128, 1, 28, 28), torch.zeros(128, 3, 32, 32)]
    assert compute_microbatch_split(x, 30) == 128 and compute_microbatch_split(x, 2000) == 80
