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


@pytest.mark.ref
def test_reference_checkpoints_load_strict_into_every_model_class(tmp_path):
    """Build container only (needs /root/reference): a checkpoint SAVED by the unmodified reference model (through
    oracle/ref_shim.py) loads with strict=True into the matching mmvae_b200 VAE_* class, and round-trips back into the
    reference (main.py:70, 240-246: `torch.save(model.state_dict())` / `load_state_dict`)."""
    import argparse
    import importlib.util
    import mmvae_b200
    from oracle import bivae_oracle as O
    from oracle import synth
    spec_ = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "make_golden.py"))
    MG = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(MG)
    ns = lambda **kw: argparse.Namespace(**{**dict(latent_dim=20, dist="laplace", recon_losses=["normal", "normal"], llik_scaling=0.0,
                                                    device="cpu", data_path="", epochs=1, beta_kl=1.0, decrease_beta_kl=1.0,
                                                    fix_jencoder=True, fix_decoders=True, no_recon=True, no_nf=False, flow="maf",
                                                    dcca=False, dim_dcca=9, num_hidden_layers=1, warmup=0), **kw})
    cases = [
        ("VAE_mnist_svhn", O.spec_mnist_svhn(20, "laplace"), MG.ref_mmvae, ns()),
        ("VAE_mmvae_msf", O.spec_msf(30, "normal"), MG.ref_mmvae, ns(latent_dim=30, dist="normal", recon_losses=["normal"] * 3)),
        ("VAE_circles_squares", O.spec_circles(2, "normal"), MG.ref_mmvae, ns(latent_dim=2, dist="normal")),
        ("VAE_mvae_mnist_svhn", O.spec_mnist_svhn(20, "normal"), MG.ref_mvae, ns(dist="normal")),
    ]
    for cls, spec, build_ref, params in cases:
        ref = build_ref(spec, synth.make_state_dict(spec, 3))
        path = tmp_path / f"{cls}.pt"
        torch.save(ref.state_dict(), path)
        ours = getattr(mmvae_b200, cls)(params)
        missing = ours.load_state_dict(torch.load(path), strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys, cls
        for k, v in ref.state_dict().items():
            assert torch.equal(ours.state_dict()[k].cpu(), v), (cls, k)
        ref.load_state_dict(ours.state_dict(), strict=True)              # and back
