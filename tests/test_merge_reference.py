"""The subgrid merge pinned on the reference's own code: tests/golden/merge_tiny.npz holds what
the unmodified ``reduce_grid_info`` / ``Q_all_memory`` / ``merge_pdf1d_stats`` / ``merge_lnp``
(tools/subgridding_tools.py:189-271, 274-539, 542-658) produced for three row slices of a grid
(oracle/make_golden.py::make_merge_case).  Checked here: the oracle's restated merge, the
package's grid-info functions and its file-level ``merge_pdf1d_stats`` / ``merge_lnp`` drop-ins
(arithmetic through a NumPy double on CPU; ``-m gpu``: through the CUDA merge kernels, with the
subgrid fits done by the CUDA path too)."""
import os

import numpy as np
import pytest

from tests import util


def load_case():
    from oracle.make_golden import unpack_inputs

    z = np.load(os.path.join(util.GOLDEN, "merge_tiny.npz"), allow_pickle=False)
    g, nz, obs = unpack_inputs(z)
    info = {str(q): {"min": z["info_min"][i], "max": z["info_max"][i],
                     "num_unique": int(z["info_num_unique"][i]),
                     "unique_vals": z["info_unique__" + str(q)]}
            for i, q in enumerate(z["info_keys"])}
    return z, g, nz, obs, info


def subgrids_of(z, g, nz):
    from beast_b200.subgrid import uniform_slices

    return [(g.subgrid(s), {k: v[s] for k, v in nz.items()}) for s in uniform_slices(len(g), int(z["n_sub"]))]


def check_merged(z, qnames, pdf1d, stats, tight=True):
    """merged pdf1d arrays + stats columns against the reference's merge"""
    rt = 1e-12 if tight else util.PDF_RTOL
    assert list(qnames) == [str(q) for q in z["ref_qnames"]]
    want_cols = [k[len("merged_stats__"):] for k in z.files if k.startswith("merged_stats__")]
    assert sorted(stats.keys()) == sorted(want_cols)  # chi2min_indx / specgrid_indx dropped, tag added
    for k, q in enumerate(qnames):
        want = z["merged_pdf1d_%03d" % k]
        np.testing.assert_array_equal(pdf1d[k][-1], want[-1])
        np.testing.assert_allclose(pdf1d[k][:-1], want[:-1], rtol=rt, atol=util.PDF_ATOL)
        np.testing.assert_array_equal(stats[q + "_Best"], z["merged_stats__" + q + "_Best"])
        # (CUDA path: parameter values in 32-bit fixed point over the parameter's range)
        span = float(np.nanmax(want[-1]) - np.nanmin(want[-1]))
        np.testing.assert_allclose(stats[q + "_Exp"], z["merged_stats__" + q + "_Exp"],
                                   rtol=1e-12 if tight else util.STAT_RTOL,
                                   atol=1e-300 if tight else 1e-9 * span + 1e-300)
        tol = util.bin_width_tolerance(want[-1])
        for p in (16, 50, 84):
            got, ref = stats["%s_p%d" % (q, p)], z["merged_stats__%s_p%d" % (q, p)]
            assert np.max(np.abs(got - ref)) <= 1e-6 * tol + 1e-9 * np.max(np.abs(ref)) + 1e-300, (q, p)
    for c in ("Pmax", "chi2min", "total_log_norm"):
        np.testing.assert_allclose(stats[c], z["merged_stats__" + c], rtol=1e-12, atol=0 if tight else util.LNP_ATOL)
    np.testing.assert_array_equal(stats["Pmax_indx"], z["merged_stats__Pmax_indx"])
    np.testing.assert_array_equal(stats["best_gridsub_tag"], z["merged_stats__best_gridsub_tag"])


def oracle_subgrid_fits(z, g, nz, obs, info):
    from oracle import fit_oracle as fo

    keys = [str(k) for k in z["keys"]]
    return [fo.fit_all(obs.fluxes, fo.OracleSetup(sg, sn, keys, max_nbins=int(z["max_nbins"]),
                                                  grid_info_dict=info, do_not_normalize=True),
                       float(z["threshold"]), fast=True) for sg, sn in subgrids_of(z, g, nz)]


def test_grid_info_matches_the_reference():
    """min / max of every quantity and the bins that follow; for the flux quantities the
    reference records the unique values of whatever column came last (subgridding_tools.py:170),
    only their count matters downstream and both counts exceed max_nbins"""
    from beast_b200 import subgrid
    from oracle import fit_oracle as fo

    z, g, nz, obs, info = load_case()
    subs = subgrids_of(z, g, nz)
    for mine in (subgrid.reduce_grid_info([subgrid.subgrid_info(a, b) for a, b in subs]),
                 fo.reduce_grid_info([fo.subgrid_info(a, b) for a, b in subs])):
        assert sorted(mine) == sorted(info)
        for q in info:
            assert mine[q]["min"] == info[q]["min"] and mine[q]["max"] == info[q]["max"], q
            if q.startswith("symlog"):
                assert min(mine[q]["num_unique"], info[q]["num_unique"]) > int(z["max_nbins"])
            else:
                assert mine[q]["num_unique"] == info[q]["num_unique"], q
                np.testing.assert_array_equal(mine[q]["unique_vals"], info[q]["unique_vals"])


def test_oracle_merge_matches_the_reference():
    from oracle import fit_oracle as fo

    z, g, nz, obs, info = load_case()
    res = oracle_subgrid_fits(z, g, nz, obs, info)
    m = fo.merge_subgrid_results(res)
    stats = {"Pmax": m["Pmax"], "chi2min": m["chi2min"], "total_log_norm": m["total_log_norm"],
             "Pmax_indx": m["Pmax_indx"], "best_gridsub_tag": m["best_gridsub_tag"]}
    for k, q in enumerate(m["qnames"]):
        stats[q + "_Best"], stats[q + "_Exp"] = m["best_vals"][:, k], m["exp_vals"][:, k]
        for i, p in enumerate((16, 50, 84)):
            stats["%s_p%d" % (q, p)] = m["per_vals"][:, k, i]
    check_merged(z, m["qnames"], m["pdf1d"], stats)
    # sparse lnP: entries kept, their order and subgrid tags
    sparse = [[(s[0], s[1].astype(np.float32).astype(float)) for s in r["sparse"]] for r in res]
    merged = fo.merge_sparse_lnp(sparse, threshold=float(z["threshold"]))
    for e, (idx, lnp, sub) in enumerate(merged):
        want = z["merged_lnp_%03d_lnp" % e]
        n = int(np.isfinite(want).sum())
        np.testing.assert_array_equal(idx, z["merged_lnp_%03d_idx" % e][:n])
        np.testing.assert_allclose(lnp, want[:n], rtol=0, atol=1e-4)
        np.testing.assert_array_equal(sub, z["merged_lnp_%03d_subgrid" % e][:n])


def write_subgrid_files(tmp_path, z, per_sub, lnp_per_sub=None):
    """stats / pdf1d (/ lnp) files of the subgrid fits, named as run_fitting names them"""
    from beast_b200 import fit

    pdf_f, stats_f, lnp_f = [], [], []
    for gi, (qnames, pdf1d, stats) in enumerate(per_sub):
        base = str(tmp_path / ("beast_gridsub%d" % gi))
        fit.save_pdf1d(base + "_pdf1d.fits", pdf1d, qnames)
        from beast_b200 import fitsio

        fitsio.write_table_file(base + "_stats.fits", [stats, {"filternames": np.array(["A", "B"]),
                                                               "wavelengths": np.array([1.0, 2.0])}])
        pdf_f.append(base + "_pdf1d.fits")
        stats_f.append(base + "_stats.fits")
        if lnp_per_sub is not None:
            fit.write_lnp_groups(base + "_lnp.hd5", {"star_%d" % e: {"idx": i, "lnp": l}
                                                     for e, (i, l) in enumerate(lnp_per_sub[gi])})
            lnp_f.append(base + "_lnp.hd5")
    return pdf_f, stats_f, lnp_f


def reference_subgrid_outputs(z):
    """the per-subgrid files' content as the reference's own fits produced it"""
    qnames = [str(q) for q in z["ref_qnames"]]
    out = []
    for gi in range(int(z["n_sub"])):
        pdf1d = [z["sub%d_pdf1d_%03d" % (gi, k)] for k in range(len(qnames))]
        pre = "sub%d_stats__" % gi
        stats = {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}
        # column order as save_stats writes it (fit.py:94-108)
        order = []
        for q in qnames:
            order += [q + "_Best", q + "_Exp", q + "_p16", q + "_p50", q + "_p84"]
        order += ["chi2min", "chi2min_indx", "Pmax", "Pmax_indx", "specgrid_indx", "total_log_norm"]
        out.append((qnames, pdf1d, {c: stats[c] for c in order}))
    return out


def test_file_level_merges_match_the_reference(tmp_path, capsys):
    """merge_pdf1d_stats / merge_lnp over files, fed the reference's own subgrid outputs"""
    from beast_b200 import fit, fitsio, subgrid
    from oracle import fit_oracle as fo

    z, g, nz, obs, info = load_case()
    per_sub = reference_subgrid_outputs(z)
    lnps = [[(s[0], s[1].astype(np.float32).astype(float)) for s in r["sparse"]]
            for r in oracle_subgrid_fits(z, g, nz, obs, info)]
    pdf_f, stats_f, lnp_f = write_subgrid_files(tmp_path, z, per_sub, lnps)
    math = util.NumpyMergeMath([a[-1] for a in per_sub[0][1]])
    base = str(tmp_path / "merged")
    p1, s1 = subgrid.merge_pdf1d_stats(pdf_f, stats_f, output_fname_base=base, math=math)
    assert (p1, s1) == (base + "_pdf1d.fits", base + "_stats.fits")
    hd = fitsio.read_fits(p1)[1:]
    stats = fitsio.read_fits(s1)
    check_merged(z, [h["EXTNAME"] for h, _ in hd], [d for _, d in hd], stats[1][1])
    assert list(stats[2][1]["filternames"]) == ["A", "B"]  # the filter table travels along
    # a second call finds the files and leaves them alone, as the reference does
    capsys.readouterr()
    subgrid.merge_pdf1d_stats(pdf_f, stats_f, output_fname_base=base, math=math)
    assert "already merged" in capsys.readouterr().out
    with pytest.raises(AssertionError):
        subgrid.merge_pdf1d_stats(pdf_f, stats_f[:-1], re_run=True, output_fname_base=base, math=math)

    name = subgrid.merge_lnp(lnp_f, output_fname_base=base, threshold=float(z["threshold"]))
    assert name == base + "_lnp.hd5"
    groups = fit.read_lnp_groups(name)
    assert sorted(groups) == sorted("star_%d" % e for e in range(len(obs)))
    for e in range(len(obs)):
        for c in ("idx", "lnp", "subgrid"):
            got, want = groups["star_%d" % e][c], z["merged_lnp_%03d_%s" % (e, c)]
            assert got.shape == want.shape  # every star padded to the longest list
            np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
            ok = np.isfinite(want)
            np.testing.assert_allclose(got[ok], want[ok], rtol=0, atol=1e-4 if c == "lnp" else 0)


@pytest.mark.gpu
def test_cuda_fits_and_cuda_merge_match_the_reference(tmp_path):
    """three subgrids fitted by the CUDA path (do_not_normalize, shared grid_info_dict), written
    to files, merged by the CUDA merge kernels: against the reference's merged files"""
    from beast_b200 import fit, fitsio, subgrid

    z, g, nz, obs, info = load_case()
    keys = [str(k) for k in z["keys"]]
    pdf_f, stats_f = [], []
    for gi, (sg, sn) in enumerate(subgrids_of(z, g, nz)):
        base = str(tmp_path / ("beast_gridsub%d" % gi))
        fit.Q_all_memory({}, obs, sg, sn, list(keys), max_nbins=int(z["max_nbins"]),
                         stats_outname=base + "_stats.fits", pdf1d_outname=base + "_pdf1d.fits",
                         grid_info_dict=info, threshold=float(z["threshold"]), do_not_normalize=True)
        pdf_f.append(base + "_pdf1d.fits")
        stats_f.append(base + "_stats.fits")
    base = str(tmp_path / "merged")
    p1, s1 = subgrid.merge_pdf1d_stats(pdf_f, stats_f, output_fname_base=base)
    hd = fitsio.read_fits(p1)[1:]
    check_merged(z, [h["EXTNAME"] for h, _ in hd], [d for _, d in hd], fitsio.read_fits(s1)[1][1], tight=False)


def test_merge_stats_files_matches_the_reference(tmp_path, monkeypatch):
    """merge_stats_files (tools/merge_beast_stats.py:12-71): row-wise stack, the reorder tag
    taken from between "_sd" and "_stats" of each file name, filter table carried along"""
    from oracle import ref_shim

    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    from beast_b200 import fitsio, merge

    rng = np.random.default_rng(5)
    monkeypatch.chdir(tmp_path)  # (the tag is searched for in the name as given)
    names = ["proj_sd0-1_sub0_stats.fits", "proj_sd2-3_sub1_stats.fits"]
    tabs = {}
    for i, n in enumerate(names):
        tabs[n] = {"Name": np.array(["s%d_%d" % (i, j) for j in range(3 + i)]),
                   "Av_Exp": rng.random(3 + i), "Pmax_indx": rng.integers(0, 99, 3 + i)}
    filt = {"filternames": np.array(["F1", "F2"]), "wavelengths": np.array([1.0, 2.0])}
    n_ref, hdus = ref_shim.run_reference_merge_stats(tabs, filt)
    for n in names:
        fitsio.write_table_file(n, [tabs[n], filt])
    out = "all"
    assert merge.merge_stats_files(names, out) == n_ref == 7
    got = fitsio.read_fits(out + "_stats.fits")
    assert list(got[1][1].keys()) == list(hdus[0].keys())
    for c in hdus[0]:
        np.testing.assert_array_equal(got[1][1][c], hdus[0][c])
    assert set(got[1][1]["reorder_tag"]) == {"sd0-1_sub0", "sd2-3_sub1"}
    np.testing.assert_array_equal(got[2][1]["filternames"], hdus[1]["filternames"])
