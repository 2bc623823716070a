"""Pins the oracle on the unmodified reference, run in-process through oracle/ref_shim.py.
Only possible where /root/reference exists (the build container); the golden fixtures carry
the same pin to the GPU box."""
import numpy as np
import pytest

from oracle import fit_oracle as fo
from oracle import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(),
                                reason="reference tree not present on this machine")


@pytest.mark.parametrize("thr,nbins,p2d,nonorm", [(-10.0, 50, ["Av", "M_ini", "logT"], False),
                                                 (-40.0, 200, None, True)])
def test_oracle_equals_reference(thr, nbins, p2d, nonorm):
    from beast_b200 import synth

    g = synth.make_sedgrid("tiny", seed=91)
    g["weight"][::9] = 0.0
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 10, seed=92)
    keys = fo.default_keys(g)
    ref = ref_shim.run_reference_fit(obs, g, nz, keys, threshold=thr, max_nbins=nbins,
                                     pdf2d_param_list=p2d, do_not_normalize=nonorm)
    st = fo.OracleSetup(g, nz, keys, max_nbins=nbins, pdf2d_param_list=p2d,
                        make_pdf2d=p2d is not None, do_not_normalize=nonorm)
    out = fo.fit_all(obs.fluxes, st, thr, fast=False)
    s = ref["stats"]
    assert s["qnames"] == out["qnames"]
    # faithful mode repeats the reference's operation sequence: identical bits
    for k in ("best_vals", "exp_vals", "per_vals", "chi2min", "chi2min_indx", "Pmax", "Pmax_indx",
              "specgrid_indx", "total_log_norm"):
        np.testing.assert_array_equal(out[k], s[k], err_msg=k)
    for a, b in zip(out["pdf1d"], ref["pdf1d"]):
        np.testing.assert_array_equal(a, b)
    for a, b in zip(out["pdf2d"], ref.get("pdf2d", [])):
        np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-16)
    for e in range(len(obs)):
        idx, lnp, chi2 = ref_shim.reference_sparse_set(obs.fluxes[e], g, nz, thr, nonorm)
        np.testing.assert_array_equal(out["sparse"][e][0], idx)


def test_mutates_caller_key_list_like_reference():
    # fit.py:456-459 appends the symlog names to the caller's list
    from beast_b200 import synth

    g = synth.make_sedgrid("tiny", seed=3)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 2, seed=4)
    fit = ref_shim.load_reference_fit()
    keys = ["Av", "M_ini"]
    import contextlib
    import io

    saved = fit.save_stats
    fit.save_stats = lambda *a, **k: None
    try:
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            fit.Q_all_memory({}, obs, g, nz, keys, stats_outname="x", threshold=-10.0)
    finally:
        fit.save_stats = saved
    assert keys[2:] == ["symlog" + f + "_wd_bias" for f in g.filters]
