"""The oracle (oracle/fit_oracle.py) against the golden vectors produced by the reference
itself (oracle/make_golden.py), plus the reference's own known answers on this path."""
import numpy as np
import pytest

from oracle import fit_oracle as fo
from tests import util


@pytest.mark.parametrize("case", util.GOLDEN_CASES)
@pytest.mark.parametrize("fast", [True, False])
def test_oracle_matches_reference_outputs(case, fast):
    g, nz, obs, args, ref = util.load_golden(case)
    st = fo.OracleSetup(g, nz, args["keys"], max_nbins=args["max_nbins"],
                        pdf2d_param_list=args["pdf2d_param_list"],
                        make_pdf2d=args["pdf2d_param_list"] is not None,
                        do_not_normalize=args["do_not_normalize"])
    out = fo.fit_all(obs.fluxes, st, args["threshold"], fast=fast)
    assert out["qnames"] == ref["qnames"]
    for k in ("chi2min_indx", "Pmax_indx", "specgrid_indx", "best_vals"):
        np.testing.assert_array_equal(out[k], ref[k])
    for k in ("Pmax", "total_log_norm", "chi2min", "exp_vals", "per_vals"):
        np.testing.assert_allclose(out[k], ref[k], rtol=1e-11, atol=1e-13)
    for a, b in zip(out["pdf1d"], ref["pdf1d"]):
        np.testing.assert_allclose(a, b, rtol=1e-11, atol=1e-15)
    assert out["pdf2d_pairs"] == ref["pdf2d_pairs"]
    for a, b in zip(out["pdf2d"], ref["pdf2d"]):
        np.testing.assert_allclose(a, b, rtol=1e-11, atol=1e-15)
    off = ref["sparse_off"]
    for e, (idx, lnp, chi2) in enumerate(out["sparse"]):
        np.testing.assert_array_equal(idx, ref["sparse_idx"][off[e]:off[e + 1]])
        np.testing.assert_allclose(lnp, ref["sparse_lnp"][off[e]:off[e + 1]], rtol=0, atol=1e-10)
        np.testing.assert_allclose(chi2, ref["sparse_chi2"][off[e]:off[e + 1]], rtol=1e-12)


def test_symlog_known_answers():
    # beast/tools/tests/test_symlog.py:8-20
    y = np.array([-1e-10, -1e-19, -1e-20, -1e-24, 0.0, 1e-24, 1e-20, 1e-19, 1e-10])
    want = [-1.00000000e01, -1.04139269e00, -3.01029996e-01, -4.34272769e-05, 0.0,
            4.34272769e-05, 3.01029996e-01, 1.04139269e00, 1.00000000e01]
    assert np.allclose(fo.symlog(y), want)


def test_percentile_with_empty_bins():
    # empty bins are knots too
    got = fo.percentile(np.arange(8.0), [16, 50, 84], np.array([0, 0, .1, 0, 0, .6, .3, 0]))
    np.testing.assert_allclose(got, [4.2, 5.2222222222, 5.9777777778], rtol=1e-9)


def test_uniform_slices_cover_grid():
    # subgridding_tools.py:21-36
    sl = fo.uniform_slices(103, 8)
    assert sl[0].start == 0 and sl[-1].stop == 103
    assert all(a.stop == b.start for a, b in zip(sl[:-1], sl[1:]))
    assert sorted({s.stop - s.start for s in sl}) == [12, 13]


def test_zero_completeness_raises():
    g, nz, obs, args, ref = util.load_golden("tiny_thr40_nonorm")
    nz = dict(nz)
    nz["completeness"] = nz["completeness"].copy()
    nz["completeness"][3, 1] = 0.0
    with pytest.raises(ValueError, match="zero completeness"):
        fo.OracleSetup(g, nz, args["keys"])
