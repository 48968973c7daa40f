"""mmvae_b200 — B200-native (sm_100a) implementation of bivae's multimodal-ELBO hot path behind the reference's
model / objective API.  Importing the package loads libmmvae_b200.so and fails loudly when it is missing."""
from . import _lib

_lib.load()

from . import objectives, ops  # noqa: E402
from .models import (MMVAE, MNIST_SVHN, CIRCLES_SQUARES, MNIST_SVHN_FASHION, Multi_VAES, VAE_circles_squares,  # noqa: E402
                     VAE_mmvae_msf, VAE_mnist_svhn)

from . import flows  # noqa: E402
from .jmvae_nf import JMVAE_NF, JMVAE_NF_DCCA_MNIST_SVHN, VAE_jnf_mnist_svhn_dcca  # noqa: E402

from .mvae import (MOEPOE, MOEPOE_MNIST_SVHN, MVAE, MVAE_MNIST_SVHN, MVAE_MNIST_SVHN_FASHION, VAE_moepoe_mnist_svhn,  # noqa: E402
                   VAE_mvae_mnist_svhn, VAE_mvae_msf)

from .mmvae_nf import MMVAE_NF, MMVAE_NF_MNIST_SVHN, VAE_mmvae_nf_mnist_svhn  # noqa: E402
from . import poe_subset  # noqa: E402

__all__ = ["MMVAE_NF", "MMVAE_NF_MNIST_SVHN", "VAE_mmvae_nf_mnist_svhn", "poe_subset", "MOEPOE", "MOEPOE_MNIST_SVHN", "VAE_moepoe_mnist_svhn", "MVAE", "MVAE_MNIST_SVHN", "MVAE_MNIST_SVHN_FASHION", "VAE_mvae_mnist_svhn", "VAE_mvae_msf", "flows", "JMVAE_NF", "JMVAE_NF_DCCA_MNIST_SVHN", "VAE_jnf_mnist_svhn_dcca","objectives", "ops", "MMVAE", "Multi_VAES", "MNIST_SVHN", "CIRCLES_SQUARES", "MNIST_SVHN_FASHION",
           "VAE_mnist_svhn", "VAE_circles_squares", "VAE_mmvae_msf"]
