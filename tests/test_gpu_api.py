"""The reference-facing API end to end on the GPU: same call as beast.fitting.fit, files on disk
in the reference's layouts, compared with the reference's own outputs (golden fixtures)."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _run(tmp_path, case, **extra):
    from beast_b200 import fit

    g, nz, obs, args, ref = util.load_golden(case)
    names = dict(stats_outname=str(tmp_path / "stats.fits"), pdf1d_outname=str(tmp_path / "pdf1d.fits"),
                 lnp_outname=str(tmp_path / "lnp.hd5"))
    if args["pdf2d_param_list"]:
        names["pdf2d_outname"] = str(tmp_path / "pdf2d.fits")
    with pytest.warns(UserWarning):
        out = fit.summary_table_memory(obs, nz, g, keys=list(args["keys"]), threshold=args["threshold"],
                                       max_nbins=args["max_nbins"], pdf2d_param_list=args["pdf2d_param_list"],
                                       do_not_normalize=args["do_not_normalize"], **names, **extra)
    return g, obs, ref, names, out


def test_summary_table_memory_writes_reference_layouts(tmp_path):
    from beast_b200 import fitsio

    g, obs, ref, names, out = _run(tmp_path, "tiny_thr10_pdf2d")
    hd = fitsio.read_fits(names["stats_outname"])
    tab = hd[1][1]
    qn = ref["qnames"]
    assert list(hd[2][1]["filternames"]) == list(g.filters)
    np.testing.assert_array_equal(tab["Pmax_indx"], ref["Pmax_indx"].astype(int))
    np.testing.assert_array_equal(tab["chi2min_indx"], ref["chi2min_indx"].astype(int))
    np.testing.assert_array_equal(tab["specgrid_indx"], ref["specgrid_indx"].astype(int))
    np.testing.assert_allclose(tab["Pmax"], ref["Pmax"], rtol=0, atol=util.LNP_ATOL)
    np.testing.assert_allclose(tab["total_log_norm"], ref["total_log_norm"], rtol=0, atol=util.LNP_ATOL)
    for k, q in enumerate(qn):
        np.testing.assert_array_equal(tab[q + "_Best"], ref["best_vals"][:, k])
        for i, p in enumerate((16, 50, 84)):
            np.testing.assert_allclose(tab["%s_p%d" % (q, p)], ref["per_vals"][:, k, i], rtol=1e-7, atol=1e-12)
    assert list(tab)[0] == "Name" and tab["Name"][0] == "noname"
    for f in g.filters:
        np.testing.assert_array_equal(tab[f], obs.fluxes[:, g.filters.index(f)])
    ph = fitsio.read_fits(names["pdf1d_outname"])
    assert ph[0][1].shape == (2, 2) and len(ph) == len(qn) + 1
    for k, q in enumerate(qn):
        assert ph[k + 1][0]["EXTNAME"] == q
        np.testing.assert_array_equal(ph[k + 1][1][-1], ref["pdf1d"][k][-1])  # bin centres
        np.testing.assert_allclose(ph[k + 1][1][:-1], ref["pdf1d"][k][:-1], rtol=util.PDF_RTOL, atol=util.PDF_ATOL)
    p2 = fitsio.read_fits(names["pdf2d_outname"])
    assert [h[0]["EXTNAME"] for h in p2[1:]] == ref["pdf2d_pairs"]
    for k in range(len(ref["pdf2d"])):
        assert p2[k + 1][1].shape == ref["pdf2d"][k].shape
        np.testing.assert_allclose(p2[k + 1][1], ref["pdf2d"][k], rtol=util.PDF_RTOL, atol=util.PDF_ATOL)
    # sparse lnP records of every star (lnp_npts=None -> the full sparse set, fit.py:699-700)
    assert len(out.save_lnp_vals) == 0  # flushed to the file
    import os

    assert os.path.isfile(names["lnp_outname"]) or os.path.isfile(names["lnp_outname"] + ".npz")


def test_q_all_memory_extends_callers_key_list_and_resumes(tmp_path):
    from beast_b200 import fit, fitsio

    g, nz, obs, args, ref = util.load_golden("tiny_thr40_nonorm")
    keys = ["Av", "M_ini", "logT"]
    stats, pdf = str(tmp_path / "s.fits"), str(tmp_path / "p.fits")
    first = fit.Q_all_memory({}, obs, g, nz, keys, max_nbins=30, stats_outname=stats, pdf1d_outname=pdf,
                             threshold=-10.0)
    assert keys[3:] == ["symlog" + f + "_wd_bias" for f in g.filters]  # fit.py:456-459
    # pretend the job died after 5 stars: later rows zero, like a partially written run
    tab = fitsio.read_fits(stats)
    cols = {k: np.array(v) for k, v in tab[1][1].items()}
    for k, v in cols.items():
        if v.dtype.kind in "fi":
            v[5:] = 0
    fitsio.write_table_file(stats, [cols, tab[2][1]])
    imgs = fitsio.read_fits(pdf)
    arrs = [np.array(h[1]) for h in imgs[1:]]
    for a in arrs:
        a[5:-1] = 0.0
    fitsio.write_image_file(pdf, arrs, [h[0]["EXTNAME"] for h in imgs[1:]])
    again = fit.Q_all_memory({}, obs, g, nz, ["Av", "M_ini", "logT"], max_nbins=30, stats_outname=stats,
                             pdf1d_outname=pdf, threshold=-10.0, resume=True)
    np.testing.assert_array_equal(again.lnp_vals, first.lnp_vals)
    np.testing.assert_array_equal(again.best_vals, first.best_vals)
    for a, b in zip(again.save_pdf1d_vals, first.save_pdf1d_vals):
        np.testing.assert_array_equal(a, b)
