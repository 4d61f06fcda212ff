"""packnet_sfm -- B200-native drop-in for the reference's self-supervised photometric-loss
path (losses/multiview_photometric_loss.py, geometry/, utils/{image,depth}.py).

Only the subsystems on that path exist here; everything is backed by the sm_100a kernels of
libpsfm_b200.so through the C ABI in include/psfm.h.  Importing this package without the
built library raises ImportError -- there is no CPU or eager-PyTorch fallback.
"""
from . import _C  # noqa: F401  (fails loudly when the CUDA library is missing)

__all__ = ['geometry', 'losses', 'utils', 'ops']
