"""TEST INFRASTRUCTURE — import shim that lets the UNMODIFIED reference (`/root/reference/src/bivae`)
run on CPU in the build container, so that golden vectors can be generated from the reference itself
(`tests/golden/make_golden.py`) and the CPU restatement `oracle/bivae_oracle.py` can be pinned to it.

It is only usable where `/root/reference` exists (the build container); nothing on the GPU box may call
it.  Nothing under `mmvae_b200/` imports this file.

What it does:
  1. registers bare package objects for `bivae`, `bivae.models`, `bivae.models.mmvae`, ... so the
     star-imports in the reference's `__init__.py` files (which need torchnet/umap/matplotlib/...) never run;
  2. fabricates `MagicMock` modules for third-party packages that are absent here and for reference
     sub-packages that are OUT OF SCOPE (analysis, vis, dataloaders, dcca, modalities);
  3. supplies a functional stub of `pythae==0.0.8` (requirements.txt:11 — not vendored, not installed):
     ModelOutput, BaseEncoder/BaseDecoder, BaseAE, VAE, VAEConfig, Encoder_VAE_MLP, Decoder_AE_MLP, and
     `pythae.models.normalizing_flows` aliased to the reference's vendored copy
     (`my_pythae/models/normalizing_flows`);
  4. disables wandb.
The MLP definitions restate upstream pythae 0.0.8 `default_architectures.py`; the state-dict layout
(`layers.0.0.weight`, `embedding.weight`, `log_var.weight`) is the one `/root/reference/model2.pt` shows.
"""
import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys
import types
from collections import OrderedDict
from unittest.mock import MagicMock

REF_ROOT = os.environ.get("MMVAE_REFERENCE_ROOT", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "src")
REF_PKG = os.path.join(REF_SRC, "bivae")


def available():
    return os.path.isdir(REF_PKG)


_MOCK_TOP = {"torchnet", "umap", "matplotlib", "seaborn", "medmnist", "nltk", "gensim", "gdown",
             "imageio", "skimage", "mpl_toolkits"}
_MOCK_REF = ("bivae.analysis", "bivae.vis", "bivae.dataloaders", "bivae.datasets", "bivae.dcca",
             "bivae.models.modalities", "bivae.models.samplers", "bivae.data_utils")


class _MockLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = MagicMock(name=spec.name)
        m.__name__ = spec.name
        m.__path__ = []
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


class _MockFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path=None, target=None):
        top = name.split(".")[0]
        if top in _MOCK_TOP or any(name == p or name.startswith(p + ".") for p in _MOCK_REF):
            return importlib.machinery.ModuleSpec(name, _MockLoader(), is_package=True)
        return None


def _bare_package(name, path):
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


class _LenientModule(types.ModuleType):
    """Module whose unknown attributes resolve to MagicMock (for unused pythae names)."""

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        return MagicMock(name=f"{self.__name__}.{item}")


def _install_pythae_stub():
    import numpy as np
    import torch
    import torch.nn as nn
    from pydantic.dataclasses import dataclass
    from typing import Tuple, Union

    class ModelOutput(OrderedDict):
        """OrderedDict with attribute and integer access (joint_encoders.py:78 indexes `[0]`)."""

        def __getitem__(self, k):
            if isinstance(k, str):
                return dict(self.items())[k]
            return tuple(self.values())[k]

        def __setattr__(self, name, value):
            super().__setitem__(name, value)
            super().__setattr__(name, value)

        def __setitem__(self, key, value):
            super().__setitem__(key, value)
            super().__setattr__(key, value)

        def to_tuple(self):
            return tuple(self[k] for k in self.keys())

    class BaseEncoder(nn.Module):
        def __init__(self):
            nn.Module.__init__(self)

    class BaseDecoder(nn.Module):
        def __init__(self):
            nn.Module.__init__(self)

    @dataclass
    class BaseConfig:
        def to_dict(self):
            return dict(self.__dict__)

    @dataclass
    class BaseAEConfig(BaseConfig):
        input_dim: Union[Tuple[int, ...], None] = None
        latent_dim: int = 10
        uses_default_encoder: bool = True
        uses_default_decoder: bool = True

    @dataclass
    class VAEConfig(BaseAEConfig):
        reconstruction_loss: str = "mse"

    @dataclass
    class VAE_IAF_Config(VAEConfig):
        n_made_blocks: int = 2
        n_hidden_in_made: int = 3
        hidden_size: int = 128

    class Encoder_VAE_MLP(BaseEncoder):
        def __init__(self, args):
            BaseEncoder.__init__(self)
            self.input_dim = args.input_dim
            self.latent_dim = args.latent_dim
            layers = nn.ModuleList()
            layers.append(nn.Sequential(nn.Linear(int(np.prod(args.input_dim)), 512), nn.ReLU()))
            self.layers = layers
            self.depth = len(layers)
            self.embedding = nn.Linear(512, self.latent_dim)
            self.log_var = nn.Linear(512, self.latent_dim)

        def forward(self, x, output_layer_levels=None):
            out = x.reshape(-1, int(np.prod(self.input_dim)))
            for layer in self.layers:
                out = layer(out)
            output = ModelOutput()
            output["embedding"] = self.embedding(out)
            output["log_covariance"] = self.log_var(out)
            return output

    class Decoder_AE_MLP(BaseDecoder):
        def __init__(self, args):
            BaseDecoder.__init__(self)
            self.input_dim = args.input_dim
            layers = nn.ModuleList()
            layers.append(nn.Sequential(nn.Linear(args.latent_dim, 512), nn.ReLU()))
            layers.append(nn.Sequential(nn.Linear(512, int(np.prod(args.input_dim))), nn.Sigmoid()))
            self.layers = layers
            self.depth = len(layers)

        def forward(self, z, output_layer_levels=None):
            out = z
            for layer in self.layers:
                out = layer(out)
            output = ModelOutput()
            output["reconstruction"] = out.reshape((z.shape[0],) + tuple(self.input_dim))
            return output

    class BaseAE(nn.Module):
        def __init__(self, model_config, encoder=None, decoder=None):
            nn.Module.__init__(self)
            self.model_name = "BaseAE"
            self.input_dim = model_config.input_dim
            self.latent_dim = model_config.latent_dim
            self.model_config = model_config
            if decoder is None:
                decoder = Decoder_AE_MLP(model_config)
                self.model_config.uses_default_decoder = True
            else:
                self.model_config.uses_default_decoder = False
            self.set_decoder(decoder)
            self.device = None

        def set_encoder(self, encoder):
            self.encoder = encoder

        def set_decoder(self, decoder):
            self.decoder = decoder

    class VAE(BaseAE):
        def __init__(self, model_config, encoder=None, decoder=None):
            BaseAE.__init__(self, model_config=model_config, decoder=decoder)
            self.model_name = "VAE"
            if encoder is None:
                encoder = Encoder_VAE_MLP(model_config)
                self.model_config.uses_default_encoder = True
            else:
                self.model_config.uses_default_encoder = False
            self.set_encoder(encoder)

    def mod(name, **attrs):
        m = _LenientModule(name)
        m.__path__ = []
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    mod("pythae")
    mod("pythae.config", BaseConfig=BaseConfig)
    mod("pythae.data")
    mod("pythae.data.datasets", BaseDataset=object)
    mod("pythae.models", VAEConfig=VAEConfig, VAE_IAF_Config=VAE_IAF_Config, VAE=VAE, BaseAE=BaseAE,
        BaseAEConfig=BaseAEConfig)
    mod("pythae.models.base", BaseAE=BaseAE, BaseAEConfig=BaseAEConfig)
    mod("pythae.models.base.base_utils", ModelOutput=ModelOutput)
    mod("pythae.models.base.base_model", BaseAE=BaseAE, BaseEncoder=BaseEncoder, BaseDecoder=BaseDecoder)
    mod("pythae.models.nn", BaseEncoder=BaseEncoder, BaseDecoder=BaseDecoder)
    mod("pythae.models.nn.default_architectures", Encoder_VAE_MLP=Encoder_VAE_MLP,
        Decoder_AE_MLP=Decoder_AE_MLP)
    mod("pythae.models.nn.benchmarks")
    mod("pythae.models.nn.benchmarks.celeba")
    mod("pythae.models.nn.benchmarks.utils")
    mod("pythae.models.vae", VAE=VAE, VAEConfig=VAEConfig)
    mod("pythae.models.vae.vae_config", VAEConfig=VAEConfig)
    mod("pythae.models.vae_iaf")
    mod("pythae.models.vae_iaf.vae_iaf_config", VAE_IAF_Config=VAE_IAF_Config)
    mod("pythae.models.vae_lin_nf")
    mod("pythae.models.vae_lin_nf.vae_lin_nf_config")
    # pythae.models.normalizing_flows := the reference's vendored copy (same arithmetic)
    nf_dir = os.path.join(REF_PKG, "my_pythae", "models", "normalizing_flows")
    spec = importlib.util.spec_from_file_location(
        "pythae.models.normalizing_flows", os.path.join(nf_dir, "__init__.py"),
        submodule_search_locations=[nf_dir])
    nf = importlib.util.module_from_spec(spec)
    sys.modules["pythae.models.normalizing_flows"] = nf
    spec.loader.exec_module(nf)


_installed = False


def install():
    """Idempotent. After this, `import bivae.objectives` etc. resolve to the reference's own files."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference not found at {REF_PKG}; the shim only works in the build container")
    os.environ.setdefault("WANDB_MODE", "disabled")
    os.environ.setdefault("WANDB_SILENT", "true")
    sys.meta_path.insert(0, _MockFinder())
    for name, rel in [("bivae", ""), ("bivae.models", "models"), ("bivae.models.mmvae", "models/mmvae"),
                      ("bivae.models.jmvae_nf", "models/jmvae_nf"), ("bivae.my_pythae", "my_pythae"),
                      ("bivae.my_pythae.models", "my_pythae/models")]:
        _bare_package(name, os.path.join(REF_PKG, rel))
    _install_pythae_stub()
    import wandb
    try:
        wandb.init(mode="disabled")
    except Exception:
        pass
    # expose the four VAE wrappers under bivae.my_pythae.models like its __init__ would (minus LinNF)
    pm = sys.modules["bivae.my_pythae.models"]
    pm.my_VAE = importlib.import_module("bivae.my_pythae.models.vae.vae_model_adapted").my_VAE
    pm.laplace_VAE = importlib.import_module("bivae.my_pythae.models.laplace_vae.laplace_vae").laplace_VAE
    pm.my_VAE_IAF = importlib.import_module("bivae.my_pythae.models.vae_iaf.vae_iaf_model_adapted").my_VAE_IAF
    maf = importlib.import_module("bivae.my_pythae.models.vae_maf")
    pm.my_VAE_MAF, pm.VAE_MAF_Config = maf.my_VAE_MAF, maf.VAE_MAF_Config
    pm.my_VAE_LinNF = MagicMock(name="my_VAE_LinNF")
    _installed = True


def ref(modname):
    """Import a reference module by dotted name, e.g. ref('bivae.objectives')."""
    install()
    return importlib.import_module(modname)
