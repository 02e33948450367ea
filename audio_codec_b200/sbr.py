"""Drop-in for the reference's `sbr` module (coder/sbr.py:1-12): which scale factor
bands are left to spectral band replication.  Config-time host logic."""
import numpy as np

SBR_FACTOR = 2

__all__ = ["SBR_FACTOR", "omitted_bands", "sbr_factor"]


def omitted_bands(sfBands, factor=SBR_FACTOR):
    sbr_thresh = sfBands.upperLine[-1] // factor
    return np.where(np.asarray(sfBands.lowerLine) >= sbr_thresh)[0]


def sbr_factor():
    return SBR_FACTOR
