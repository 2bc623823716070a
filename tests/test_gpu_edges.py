"""Edge cases of the CUDA path: more than 32 parameters (two lane rounds), 8 bands, one star,
ragged star counts, stars and models the reference could not rank."""
import warnings

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _fit(g, nz, fluxes, keys, thr, nbins, **kw):
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        setup = FitSetup(g, nz, list(keys), max_nbins=nbins)
    return setup, FitEngine(setup).fit(fluxes, threshold=thr, **kw)


def test_eight_bands_38_parameters_against_oracle():
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid((5, 9, 2, 6, 3, 2, 1), synth.UVNIR_FILTERS, synth.UVNIR_LAMB, seed=51)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 37, seed=52)  # not a multiple of anything
    keys = fo.default_keys(g)
    want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=200), -40.0, fast=True)
    setup, res = _fit(g, nz, obs.fluxes, keys, -40.0, 200, sparse="all")
    assert setup.n_q == 38 and setup.nb == 8
    util.compare_to_reference(setup, res, want)
    off = np.zeros(len(obs) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(s[0]) for s in want["sparse"]])
    util.compare_sparse(setup, res, off, np.concatenate([s[0] for s in want["sparse"]]),
                        np.concatenate([s[1] for s in want["sparse"]]),
                        np.concatenate([s[2] for s in want["sparse"]]))


def test_single_star_and_single_band():
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("tiny", seed=61)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 1, seed=62)
    keys = ["Av", "M_ini", "distance"]  # distance: one unique value -> one bin
    want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=10), -10.0, fast=True)
    setup, res = _fit(g, nz, obs.fluxes, keys, -10.0, 10)
    util.compare_to_reference(setup, res, want)
    # one band only
    g1 = synth.SyntheticSEDGrid(g.lamb[:1], g.seds[:, :1], {k: g[k] for k in g.keys()}, g.filters[:1])
    nz1 = {k: v[:, :1] for k, v in nz.items()}
    f1 = obs.fluxes[:, :1]
    want1 = fo.fit_all(f1, fo.OracleSetup(g1, nz1, keys, max_nbins=10), -10.0, fast=True)
    setup1, res1 = _fit(g1, nz1, f1, keys, -10.0, 10)
    util.compare_to_reference(setup1, res1, want1)


def test_unrankable_star_and_models():
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("tiny", seed=71)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 5, seed=72)
    keys = ["Av", "logT"]
    flux = obs.fluxes.copy()
    flux[2, 3] = np.nan  # the reference raises on this star (max of an empty list, fit.py:670)
    setup, res = _fit(g, nz, flux, keys, -10.0, 30)
    assert np.isnan(res.Pmax[2]) and res.Pmax_j[2] == -1 and res.n_sparse[2] == 0
    assert np.isnan(res.total_log_norm[2]) and not res.pdf1d[2].any() and not res.per_vals[2].any()
    ok = [0, 1, 3, 4]
    want = fo.fit_all(obs.fluxes[ok], fo.OracleSetup(g, nz, keys, max_nbins=30), -10.0, fast=True)
    np.testing.assert_array_equal(setup.g0_indxs[res.Pmax_j[ok]], want["Pmax_indx"].astype(np.int64))
    np.testing.assert_allclose(res.total_log_norm[ok], want["total_log_norm"], rtol=0, atol=util.LNP_ATOL)
    # models whose noise model is not finite are dropped with a warning (deviations)
    nz2 = {k: v.copy() for k, v in nz.items()}
    nz2["error"][10, 0] = 0.0
    from beast_b200.gridprep import FitSetup

    with pytest.warns(UserWarning, match="non-finite"):
        st = FitSetup(g, nz2, list(keys), max_nbins=30)
    assert 10 not in st.g0_indxs and st.n_models == len(g) - 1


def test_high_signal_to_noise_selection_margin():
    """S/N 5000 per band: the FP32 sweep is 50x less accurate than in the benchmark grids; the
    error-bound margin must still deliver the exact sparse set and arg-max."""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("small", seed=81)
    nz = synth.make_noisemodel(g, snr_cap=5000.0, depth_scale=1e-3)
    obs = synth.make_observations(g, nz, 40, seed=82)
    keys = ["Av", "M_ini", "logA"]
    want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=50), -10.0, fast=True)
    setup, res = _fit(g, nz, obs.fluxes, keys, -10.0, 50, sparse="all")
    util.compare_to_reference(setup, res, want)
    np.testing.assert_array_equal(res.n_sparse, [len(s[0]) for s in want["sparse"]])


@pytest.mark.parametrize("bands", [6, 8])
def test_full_covariance_against_oracle(bands):
    """trunchen-like noise model: N_covar_logLikelihood path (fit.py:432-444, likelihood.py:58-106)"""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    if bands == 8:
        g = synth.make_sedgrid((5, 9, 2, 6, 3, 2, 1), synth.UVNIR_FILTERS, synth.UVNIR_LAMB, seed=91)
    else:
        g = synth.make_sedgrid("small", seed=91)
    nz = synth.make_noisemodel_fullcov(g)
    obs = synth.make_observations(g, nz, 33, seed=92)
    keys = ["Av", "Rv", "M_ini", "logA", "logT", "logL"]
    ost = fo.OracleSetup(g, nz, keys, max_nbins=60)
    assert ost.full_cov
    want = fo.fit_all(obs.fluxes, ost, -10.0, fast=True)
    setup, res = _fit(g, nz, obs.fluxes, keys, -10.0, 60, sparse="all")
    assert setup.full_cov
    util.compare_to_reference(setup, res, want)
    off = np.zeros(len(obs) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(s[0]) for s in want["sparse"]])
    util.compare_sparse(setup, res, off, np.concatenate([s[0] for s in want["sparse"]]),
                        np.concatenate([s[1] for s in want["sparse"]]),
                        np.concatenate([s[2] for s in want["sparse"]]))
