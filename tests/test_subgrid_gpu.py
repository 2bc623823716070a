"""Subgrid mode on the GPU: fit 3 row slices of a grid separately and merge them on the device;
the oracle is per-subgrid reference-equivalent fits + the restated merge arithmetic
(merged != full-grid fit by construction, so that is the right oracle)."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def test_three_subgrids_merge_matches_oracle():
    import torch

    from beast_b200 import subgrid, synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup
    from oracle import fit_oracle as fo

    g = synth.make_sedgrid("small", seed=31)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 20, seed=32)
    keys = fo.default_keys(g)
    slices = subgrid.uniform_slices(len(g), 3)
    subs = [(g.subgrid(s), {k: v[s] for k, v in nz.items()}) for s in slices]
    info = subgrid.reduce_grid_info([subgrid.subgrid_info(sg, sn) for sg, sn in subs])
    oinfo = fo.reduce_grid_info([fo.subgrid_info(sg, sn) for sg, sn in subs])
    assert sorted(info) == sorted(oinfo)
    for q in info:
        assert info[q]["min"] == oinfo[q]["min"] and info[q]["num_unique"] == oinfo[q]["num_unique"]

    thr = -40.0
    parts, oparts, setup0 = [], [], None
    for sg, sn in subs:
        st = FitSetup(sg, sn, list(keys), max_nbins=50, grid_info_dict=info, do_not_normalize=True)
        setup0 = setup0 or st
        res = FitEngine(st).fit(obs.fluxes, threshold=thr)
        parts.append(subgrid.local_result_tensors(st, res, torch.device("cuda")))
        ost = fo.OracleSetup(sg, sn, keys, max_nbins=50, grid_info_dict=oinfo, do_not_normalize=True)
        oparts.append(fo.fit_all(obs.fluxes, ost, thr, fast=True, keep_sparse=False))
    merged = subgrid.merge_subgrids(parts, subgrid.CudaMergeMath(setup0))
    want = fo.merge_subgrid_results(oparts)

    np.testing.assert_array_equal(merged["best_gridsub_tag"].cpu().numpy(), want["best_gridsub_tag"])
    np.testing.assert_array_equal(merged["Pmax_indx"].cpu().numpy(), want["Pmax_indx"])
    np.testing.assert_array_equal(merged["best_vals"].cpu().numpy(), want["best_vals"])
    np.testing.assert_allclose(merged["Pmax"].cpu().numpy(), want["Pmax"], rtol=0, atol=util.LNP_ATOL)
    np.testing.assert_allclose(merged["total_log_norm"].cpu().numpy(), want["total_log_norm"], rtol=0,
                               atol=util.LNP_ATOL)
    np.testing.assert_allclose(merged["chi2min"].cpu().numpy(), want["chi2min"], rtol=util.STAT_RTOL)
    span = np.array([c.max() - c.min() for c in setup0.qcols]) + 1.0
    assert np.all(np.abs(merged["exp_vals"].cpu().numpy() - want["exp_vals"])
                  <= util.STAT_RTOL * np.abs(want["exp_vals"]) + 1e-8 * span)
    pdf = merged["pdf1d"].cpu().numpy()
    per = merged["per_vals"].cpu().numpy()
    for k in range(setup0.n_q):
        a, b = setup0.bin_off[k], setup0.bin_off[k + 1]
        np.testing.assert_allclose(pdf[:, a:b], want["pdf1d"][k][:-1], rtol=util.PDF_RTOL, atol=util.PDF_ATOL)
        tol = util.bin_width_tolerance(setup0.bin_val[a:b])
        assert np.max(np.abs(per[:, k, :] - want["per_vals"][:, k, :])) <= 1e-6 * tol + 1e-12


def test_resident_subgrids_merged_per_batch_at_size():
    """two subgrids of 2e6 models each (C2's grid cut in half), resident on one GPU, every batch
    merged on the device by SubgridFit -- against per-subgrid oracle fits + the restated merge
    (pinned on the reference's merge_pdf1d_stats: tests/test_merge_reference.py) on six stars"""
    import warnings

    from beast_b200 import subgrid, synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup
    from oracle import fit_oracle as fo

    axes = synth.CONFIG_AXES["C2"]
    n = int(np.prod(axes))
    slices = subgrid.uniform_slices(n, 2)
    subs = []
    for sl in slices:
        g = synth.make_sedgrid("C2", seed=1234, rows=(sl.start, sl.stop))
        subs.append((g, synth.make_noisemodel(g)))
    obs = synth.make_observations(subs[0][0], subs[0][1], 48, seed=99, with_coords=False)
    keys = fo.default_keys(subs[0][0])
    info = subgrid.reduce_grid_info([subgrid.subgrid_info(g, nz) for g, nz in subs])
    engines = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for g, nz in subs:
            st = FitSetup(g, nz, list(keys), max_nbins=200, grid_info_dict=info, do_not_normalize=True)
            assert st.n_models == 2_000_000
            engines.append(FitEngine(st, scratch_budget_bytes=4 << 30))
    sf = subgrid.SubgridFit(engines, subgrid.CudaMergeMath(engines[0].setup))
    m = sf.fit_batch(obs.fluxes, -10.0)
    assert m["rows"] == (0, len(obs))
    pick = [0, 5, 11, 23, 40, 47]
    oparts = []
    for g, nz in subs:
        ost = fo.OracleSetup(g, nz, keys, max_nbins=200, grid_info_dict=info, do_not_normalize=True)
        oparts.append(fo.fit_all(obs.fluxes[pick], ost, -10.0, fast=True, keep_sparse=False))
    want = fo.merge_subgrid_results(oparts)
    got = {k: (v.cpu().numpy()[pick] if hasattr(v, "cpu") else v) for k, v in m.items()}
    np.testing.assert_array_equal(got["best_gridsub_tag"], want["best_gridsub_tag"])
    np.testing.assert_array_equal(got["Pmax_indx"], want["Pmax_indx"])
    np.testing.assert_array_equal(got["best_vals"], want["best_vals"])
    np.testing.assert_allclose(got["Pmax"], want["Pmax"], rtol=0, atol=util.LNP_ATOL)
    np.testing.assert_allclose(got["total_log_norm"], want["total_log_norm"], rtol=0, atol=util.LNP_ATOL)
    np.testing.assert_allclose(got["chi2min"], want["chi2min"], rtol=util.STAT_RTOL)
    st0 = engines[0].setup
    span = np.array([info[q]["max"] - info[q]["min"] for q in st0.qnames]) + 1e-300
    assert np.all(np.abs(got["exp_vals"] - want["exp_vals"]) <= util.STAT_RTOL * np.abs(want["exp_vals"]) + 1e-8 * span)
    for k in range(st0.n_q):
        a, b = st0.bin_off[k], st0.bin_off[k + 1]
        np.testing.assert_allclose(got["pdf1d"][:, a:b], want["pdf1d"][k][:-1], rtol=util.PDF_RTOL, atol=util.PDF_ATOL)
        tol = util.bin_width_tolerance(st0.bin_val[a:b])
        assert np.max(np.abs(got["per_vals"][:, k, :] - want["per_vals"][:, k, :])) <= 1e-6 * tol + 1e-12
