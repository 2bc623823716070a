"""BASELINE.json sizes (C2: 4,000,000 models x 6 bands) on the GPU: size-independent
properties on 1536 stars plus a direct oracle comparison on a handful of them."""
import warnings

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2():
    from beast_b200 import synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("C2", seed=1234)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 1536, seed=99, with_coords=False)
    keys = fo.default_keys(g)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        setup = FitSetup(g, nz, list(keys), max_nbins=200)
    eng = FitEngine(setup, sparse_pass="both")
    res = eng.fit(obs.fluxes, threshold=-10.0, sparse=500, lnp_rng="device", seed=7)
    return g, nz, obs, keys, setup, eng, res


def test_full_size_properties(c2):
    g, nz, obs, keys, setup, eng, res = c2
    ns = len(obs)
    assert setup.n_models == 4_000_000 and setup.n_q == 32
    assert np.all(res.Pmax_j >= 0) and np.all(res.Pmax_j < setup.n_models)
    assert np.all(res.n_sparse >= 1)
    # total_log_norm = Pmax + ln(sum of weights), weights in (e^-10, 1], the best model has weight 1
    dlt = res.total_log_norm - res.Pmax
    assert np.all(dlt >= -1e-12) and np.all(dlt <= np.log(res.n_sparse.astype(float)) + 1e-9)
    # every parameter's pdf is a normalised histogram of the same weights
    for k in range(setup.n_q):
        a, b = setup.bin_off[k], setup.bin_off[k + 1]
        row = res.pdf1d[:, a:b]
        assert np.all(row >= 0)
        np.testing.assert_allclose(row.sum(axis=1), 1.0, rtol=0, atol=1e-9)
        # percentiles are ordered and inside the bin range
        per = res.per_vals[:, k, :]
        assert np.all(np.diff(per, axis=1) >= -1e-12 * (1 + np.abs(per[:, :2])))
        lo, hi = setup.bin_val[a:b].min(), setup.bin_val[a:b].max()
        assert np.all(per >= lo - 1e-9 * abs(lo)) and np.all(per <= hi + 1e-9 * abs(hi))
        # the expectation lies inside the parameter's range
        col = setup.qcols[k][setup.g0_indxs]
        span = col.max() - col.min()
        assert np.all(res.exp_vals[:, k] >= col.min() - 1e-9 * span - 1e-300)
        assert np.all(res.exp_vals[:, k] <= col.max() + 1e-9 * span + 1e-300)
    # chi2min belongs to a member: it cannot beat the best chi2 of the arg-max by construction only
    assert np.all(res.chi2min >= 0) and np.all(res.chi2min_j >= 0)
    # sparse records: at most lnp_npts per star, distinct, lnp within the threshold of Pmax
    cnt = np.diff(res.sparse_off)
    np.testing.assert_array_equal(cnt, np.minimum(res.n_sparse, 500))
    star = np.repeat(np.arange(ns), cnt)
    assert np.all(res.sparse_lnp - res.Pmax[star].astype(np.float32) <= 1e-3)
    assert np.all(res.sparse_lnp - res.Pmax[star].astype(np.float32) > -10.0 - 1e-3)
    for e in range(0, ns, 97):
        seg = res.sparse_j[res.sparse_off[e]:res.sparse_off[e + 1]]
        assert len(np.unique(seg)) == len(seg)
    # a second run is bit-identical (fixed summation order everywhere)
    again = eng.fit(obs.fluxes, threshold=-10.0, sparse=500, lnp_rng="device", seed=7)
    for name in ("Pmax", "total_log_norm", "pdf1d", "exp_vals", "per_vals", "sparse_j"):
        np.testing.assert_array_equal(getattr(res, name), getattr(again, name), err_msg=name)


def test_full_size_against_oracle_on_a_few_stars(c2):
    from oracle import fit_oracle as fo

    g, nz, obs, keys, setup, eng, res = c2
    pick = [0, 1, 511, 777, 1200, 1535]
    ost = fo.OracleSetup(g, nz, keys, max_nbins=200)
    want = fo.fit_all(obs.fluxes[pick], ost, -10.0, fast=True)

    class Sub:
        pass

    sub = Sub()
    for name in ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "exp_vals", "pdf1d", "per_vals"):
        setattr(sub, name, getattr(res, name)[pick])
    util.compare_to_reference(setup, sub, want)
    np.testing.assert_array_equal(res.n_sparse[pick], [len(s[0]) for s in want["sparse"]])


def test_full_size_histogram_kernels_agree(c2, monkeypatch):
    """C2 size, 200 bins: the full-width histogram kernel (+ expectation kernel) and the
    windowed kernel (ranges, four-at-a-time updates, in-kernel expectation sums for broad stars)
    are two independent routes to the same sums; everything else is shared and must be equal."""
    g, nz, obs, keys, setup, eng, res = c2
    monkeypatch.setenv("BF_HIST", "full")
    alt = eng.fit(obs.fluxes, threshold=-10.0, sparse=500, lnp_rng="device", seed=7)
    for name in ("Pmax", "Pmax_j", "chi2min", "chi2min_j", "n_sparse", "sparse_j", "sparse_lnp"):
        np.testing.assert_array_equal(getattr(res, name), getattr(alt, name), err_msg=name)
    # different summation orders of the same positive terms (block sizes 128 and 512)
    np.testing.assert_allclose(res.total_log_norm, alt.total_log_norm, rtol=1e-14, atol=0)
    np.testing.assert_allclose(res.pdf1d, alt.pdf1d, rtol=1e-11, atol=1e-300)
    for k in range(setup.n_q):
        col = setup.qcols[k][setup.g0_indxs]
        span = float(col.max() - col.min())
        np.testing.assert_allclose(res.exp_vals[:, k], alt.exp_vals[:, k], rtol=1e-12,
                                   atol=1e-9 * span + 1e-300, err_msg=f"Exp {k}")
        a, b = setup.bin_off[k], setup.bin_off[k + 1]
        # np.interp over a CDF with flat stretches (empty bins) can move a percentile across the
        # stretch on a last-bit change of the pdf: within one bin always, equal almost always
        width = util.bin_width_tolerance(setup.bin_val[a:b])
        diff = np.abs(res.per_vals[:, k, :] - alt.per_vals[:, k, :])
        assert np.all(diff <= width), k
        assert np.mean(diff > 1e-6 * width) < 2e-3, (k, float(np.mean(diff > 1e-6 * width)))


def test_full_size_split_and_legacy_kernels_agree(c2):
    """C2 size: group-ordered weights + class-split histograms (the default) against star-ordered
    weights + the windowed histogram kernel"""
    g, nz, obs, keys, setup, eng, res = c2
    assert eng.split
    eng.force_legacy = True
    try:
        alt = eng.fit(obs.fluxes, threshold=-10.0, sparse=500, lnp_rng="device", seed=7)
    finally:
        eng.force_legacy = False
    for name in ("Pmax", "Pmax_j", "chi2min", "chi2min_j", "n_sparse", "sparse_j", "sparse_lnp"):
        np.testing.assert_array_equal(getattr(res, name), getattr(alt, name), err_msg=name)
    np.testing.assert_allclose(res.total_log_norm, alt.total_log_norm, rtol=1e-14, atol=0)
    np.testing.assert_allclose(res.pdf1d, alt.pdf1d, rtol=1e-11, atol=1e-300)
    for k in range(setup.n_q):
        col = setup.qcols[k][setup.g0_indxs]
        span = float(col.max() - col.min())
        np.testing.assert_allclose(res.exp_vals[:, k], alt.exp_vals[:, k], rtol=1e-12,
                                   atol=1e-9 * span + 1e-300, err_msg=f"Exp {k}")
        a, b = setup.bin_off[k], setup.bin_off[k + 1]
        width = util.bin_width_tolerance(setup.bin_val[a:b])
        assert np.all(np.abs(res.per_vals[:, k, :] - alt.per_vals[:, k, :]) <= width), k


def test_c4_size_with_pdf2d_over_several_ranges():
    """BASELINE.json configs[3] shape: 1e7 models x 8 bands, 38 parameters, 2D PDFs of three
    parameters; the scratch is kept small so that the sparse pass (and bf_pdf2d_from_weights with
    it) runs over several ranges of stars.  Six stars against the oracle, all 48 for consistency."""
    from beast_b200 import synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("C4", synth.UVNIR_FILTERS, synth.UVNIR_LAMB, seed=1234)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 48, seed=77, with_coords=False)
    keys = fo.default_keys(g)
    p2d = ["Av", "M_ini", "logT"]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        setup = FitSetup(g, nz, list(keys), max_nbins=200, pdf2d_param_list=p2d, make_pdf2d=True)
    assert setup.n_models == 10_000_000 and setup.nb == 8 and setup.n_q == 38 and len(setup.pairs) == 3
    eng = FitEngine(setup, scratch_budget_bytes=1 << 30)
    eng.max_range_entries = 30000  # (48 stars list about 110,000 groups)
    assert eng.split and setup.fast_pad == 32  # 19 fast parameters: one chunk per instruction
    res = eng.fit(obs.fluxes, threshold=-10.0, want_pdf2d=True, sparse="all")
    assert eng.last_state[1].n_ranges >= 3, "the lists fit one range: lower the scratch budget"
    pick = [0, 7, 13, 21, 34, 47]
    ost = fo.OracleSetup(g, nz, keys, max_nbins=200, pdf2d_param_list=p2d, make_pdf2d=True)
    want = fo.fit_all(obs.fluxes[pick], ost, -10.0, fast=True)

    class Sub:
        pass

    sub = Sub()
    for name in ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "exp_vals", "pdf1d", "per_vals",
                 "pdf2d"):
        setattr(sub, name, getattr(res, name)[pick])
    util.compare_to_reference(setup, sub, want, n_pdf2d=3)
    np.testing.assert_array_equal(res.n_sparse[pick], [len(s[0]) for s in want["sparse"]])
    for i, e in enumerate(pick):
        a, b = res.sparse_off[e], res.sparse_off[e + 1]
        np.testing.assert_array_equal(setup.g0_indxs[res.sparse_j[a:b]], want["sparse"][i][0])
    # every 2D PDF is a normalised histogram of the same weights as the 1D ones
    for k, pr in enumerate(setup.pairs):
        blk = res.pdf2d[:, setup.pair_off[k]:setup.pair_off[k + 1]]
        np.testing.assert_allclose(blk.sum(axis=1), 1.0, rtol=0, atol=1e-9)
