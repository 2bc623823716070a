"""
TEST INFRASTRUCTURE -- NumPy restatement of the index logic of
``beast.fitting.trim_grid.trim_models`` (reference ``beast/fitting/trim_grid.py:53-118``).
Only tests / bench may import this.  Pinned on the reference itself through
``oracle/ref_shim.run_reference_trim`` (tests/test_oracle_vs_reference.py) and the golden
fixture ``tests/golden/trim_small.npz``.
"""
import numpy as np


def data_range(fluxes):
    """faintest / brightest observed flux per band (trim_grid.py:54-72, inFlux=True)"""
    return np.amin(fluxes, axis=0), np.amax(fluxes, axis=0)


def trim_indices(seds, bias, unc, completeness, min_data, max_data, sigma_fac=3.0):
    """ascending indices of the models kept (trim_grid.py:84-118)"""
    nb = seds.shape[1]
    (indxs,) = np.where(np.sum(completeness > 0, axis=1) >= nb)
    if len(indxs) <= 0:
        raise ValueError("no models are brighter than the minimum ASTs run")
    for k in range(nb):
        val = seds[indxs, k] + bias[indxs, k]
        down = val - sigma_fac * unc[indxs, k]
        up = val + sigma_fac * unc[indxs, k]
        (sel,) = np.where((up >= min_data[k]) & (down <= max_data[k]))
        if len(sel) > 0:  # a band that would leave nothing is skipped
            indxs = indxs[sel]
    if len(indxs) == 0:
        raise ValueError("no models that are within the data range")
    return indxs
