"""trim_models (the step before the fit, reference fitting/trim_grid.py): oracle vs the
reference's own output, host decision logic, and the CUDA path vs the oracle."""
import os

import numpy as np
import pytest

from oracle import ref_shim, trim_oracle
from tests.util import GOLDEN


def _golden():
    z = np.load(os.path.join(GOLDEN, "trim_tiny.npz"))
    nz = {"bias": z["bias"], "error": z["error"], "completeness": z["completeness"]}
    return z, nz


def test_trim_oracle_matches_reference_golden():
    z, nz = _golden()
    lo, hi = trim_oracle.data_range(z["obs_flux"])
    idx = trim_oracle.trim_indices(z["seds"], nz["bias"], nz["error"], nz["completeness"], lo, hi)
    np.testing.assert_array_equal(idx, z["ref_indxs"])
    np.testing.assert_array_equal(z["seds"][idx], z["ref_seds"])
    assert 0 < len(idx) < len(z["seds"])


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not present")
def test_trim_oracle_matches_reference_run():
    from beast_b200 import synth

    g = synth.make_sedgrid("small", seed=5)
    nz = synth.make_noisemodel(g)
    nz["completeness"][::11, 2] = 0.0
    obs = synth.make_observations(g, nz, 40, seed=6)
    obs = synth.SyntheticObservations(obs.fluxes[np.argsort(obs.fluxes[:, 3])[20:]], g.filters)
    cap = ref_shim.run_reference_trim(g, nz, obs)
    lo, hi = trim_oracle.data_range(obs.fluxes)
    idx = trim_oracle.trim_indices(g.seds, nz["bias"], nz["error"], nz["completeness"], lo, hi)
    np.testing.assert_array_equal(idx, cap["cols"]["fullgrid_idx"])


def _patterns(seds, nz, lo, hi, f=3.0):
    nb = seds.shape[1]
    val = seds + nz["bias"]
    ok = ((val + f * nz["error"]) >= lo) & ((val - f * nz["error"]) <= hi)
    pat = (ok * (1 << np.arange(nb))).sum(axis=1) | ((nz["completeness"] > 0).all(axis=1).astype(np.int64) << nb)
    return np.bincount(pat, minlength=1 << (nb + 1)), pat


def test_resolve_required_reproduces_skip_rule():
    from beast_b200.trim import COMPLETE_BIT, resolve_required

    z, nz = _golden()
    lo, hi = trim_oracle.data_range(z["obs_flux"])
    lo = lo.copy()
    lo[4] = 1e30  # no model can reach this band's faintest datum: the reference skips the band
    hist, pat = _patterns(z["seds"], nz, lo, hi)
    nb = z["seds"].shape[1]
    req, n_ast, n_keep = resolve_required(hist, nb)
    assert not req & (1 << 4) and req & COMPLETE_BIT
    want = trim_oracle.trim_indices(z["seds"], nz["bias"], nz["error"], nz["completeness"], lo, hi)
    assert n_keep == len(want) and n_ast == int((nz["completeness"] > 0).all(axis=1).sum())
    mask = (req & 0xFFFF) | (1 << nb)
    np.testing.assert_array_equal(np.where((pat & mask) == mask)[0], want)
    with pytest.raises(ValueError, match="minimum ASTs"):
        resolve_required(np.bincount([3], minlength=1 << (nb + 1)), nb)


@pytest.mark.gpu
@pytest.mark.parametrize("skip_band", [False, True])
def test_gpu_trim_matches_oracle(skip_band):
    from beast_b200 import synth, trim

    z, nz = _golden()
    lo, hi = trim_oracle.data_range(z["obs_flux"])
    if skip_band:
        lo = lo.copy()
        lo[2] = 1e30
    got, n_ast = trim.trim_indices(z["seds"], nz, lo, hi)
    want = trim_oracle.trim_indices(z["seds"], nz["bias"], nz["error"], nz["completeness"], lo, hi)
    np.testing.assert_array_equal(got, want)
    if not skip_band:
        np.testing.assert_array_equal(got, z["ref_indxs"])
    # the drop-in call on a larger grid, then fit the trimmed grid
    g = synth.make_sedgrid("small", seed=5)
    nzs = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nzs, 40, seed=6)
    obs = synth.SyntheticObservations(obs.fluxes[np.argsort(obs.fluxes[:, 3])[20:]], g.filters)
    cols, tseds, tnoise, indxs = trim.trim_models(g, nzs, obs)
    lo2, hi2 = trim_oracle.data_range(obs.fluxes)
    np.testing.assert_array_equal(
        indxs, trim_oracle.trim_indices(g.seds, nzs["bias"], nzs["error"], nzs["completeness"], lo2, hi2))
    np.testing.assert_array_equal(cols["fullgrid_idx"], indxs)
    np.testing.assert_array_equal(tseds, g.seds[indxs])
    assert set(tnoise) == {"bias", "error", "completeness"} and tnoise["error"].shape == tseds.shape
