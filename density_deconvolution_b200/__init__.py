"""B200-native deconvolution-GMM hot path with the reference's ``deconv.gmm`` estimator API.

``from density_deconvolution_b200 import DeconvGMM, OnlineDeconvGMM, BatchDeconvGMM,
SGDDeconvGMM`` mirrors ``deconv.gmm`` of bayesiains/density-deconvolution; the maths runs
in hand-written sm_100a CUDA kernels behind the C ABI of ``include/dgmm_b200.h``.
"""
from ._lib import DgmmError, LIB_PATH, build  # noqa: F401

__all__ = ["DgmmError", "LIB_PATH", "build"]


def __getattr__(name):
    # estimator classes import torch; keep `import density_deconvolution_b200` light
    if name in ("DeconvGMM", "OnlineDeconvGMM", "BatchDeconvGMM", "SGDDeconvGMM", "SGDDeconvGMMModule",
                "SGDDeconvDataset", "DeconvDataset", "SGDGMM", "SGDGMMModule", "BaseSGDGMM",
                "BaseSGDGMMTim", "SGDGMMTim", "SGDDeconvGMMTim"):
        from . import estimators
        return getattr(estimators, name)
    raise AttributeError(name)
