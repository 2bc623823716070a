"""Host-side mirror of the reference interface: setup, layouts, errors, writers (CPU only)."""
import numpy as np
import pytest

from beast_b200 import fitsio, synth
from beast_b200.gridprep import FitSetup
from oracle import fit_oracle as fo
from tests import util


@pytest.fixture(scope="module")
def small():
    g = synth.make_sedgrid("tiny", seed=21)
    g["weight"][::5] = 0.0
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 6, seed=22)
    return g, nz, obs


def test_setup_matches_oracle_setup(small):
    g, nz, obs = small
    keys = fo.default_keys(g)
    with pytest.warns(UserWarning, match="zero weight"):
        st = FitSetup(g, nz, list(keys), max_nbins=50, pdf2d_param_list=["Av", "logT"], make_pdf2d=True)
    ost = fo.OracleSetup(g, nz, keys, max_nbins=50, pdf2d_param_list=["Av", "logT"], make_pdf2d=True)
    assert st.qnames == ost.qnames
    np.testing.assert_array_equal(st.g0_indxs, ost.g0_indxs)
    nm = st.n_models
    np.testing.assert_array_equal(st.lnw[:nm], ost.g0_weights)
    np.testing.assert_array_equal(st.msb[:, :nm], ost.model_wbias[ost.g0_indxs].T)
    np.testing.assert_array_equal(st.ivar[:, :nm], ost.ivar[ost.g0_indxs].T)
    np.testing.assert_array_equal(st.lnq[:nm], ost.lnq[ost.g0_indxs])
    for k, (b, ob) in enumerate(zip(st.bins, ost.bins)):
        np.testing.assert_array_equal(b.bin_vals, ob.bin_vals)
        np.testing.assert_array_equal(b.bin_num, ob.bin_num)
    assert [p.name for p in st.pairs] == ost.pairs
    for p, op in zip(st.pairs, ost.bins2d):
        np.testing.assert_array_equal(p.bin_num, op.bin_num)


def test_device_layouts(small):
    g, nz, obs = small
    st = FitSetup(g, nz, ["Av", "M_ini", "logT"], max_nbins=20)
    nb, nm = st.nb, st.n_models
    assert st.ld_model % 256 == 0 and st.tiles.shape == (st.n_tiles, 2 * nb + 1, 256)
    rows = st.tiles.transpose(1, 0, 2).reshape(2 * nb + 1, -1)
    # acc = sum (f*r - a)^2 + c2 reproduces -2 (lnP - c_off) of the oracle
    ost = fo.OracleSetup(g, nz, ["Av", "M_ini", "logT"], max_nbins=20)
    f = obs.fluxes[0]
    lnp, _ = fo.lnp_chi2_nm(f, ost.model_wbias, ost.ivar, lnq=ost.lnq)
    lnp = lnp[ost.g0_indxs] + ost.g0_weights
    fs = (f / st.band_scale).astype(np.float32).astype(np.float64)
    acc = rows[2 * nb, :nm].astype(np.float64)
    for b in range(nb):
        acc += (fs[b] * rows[b, :nm] + rows[nb + b, :nm]) ** 2
    np.testing.assert_allclose(-0.5 * acc + st.c_off, lnp, rtol=0, atol=5e-3)
    assert np.all(rows[2 * nb, nm:] > 1e38) and rows[2 * nb, :nm].min() == 0.0
    # chunk-major tables
    assert st.bin_tab.shape == (st.ld_model // 8, st.n_q, 8) and st.bin_tab.dtype == np.uint8
    j = 37
    for k in range(st.n_q):
        assert st.bin_tab[j // 8, k, j % 8] == st.bins[k].bin_num[st.g0_indxs[j]]
        dec = st.q_lin[k, 0] + st.q_lin[k, 1] * float(st.q_tab[j // 8, k, j % 8])
        col = st.qcols[k][st.g0_indxs]
        assert abs(dec - col[j]) <= (col.max() - col.min()) * 2.0**-32 + 1e-300
    assert st.err_a > 0 and st.err_b > 0


def test_group_boxes_bound_the_accumulator(small):
    """the boxes of the pruned dense pass: for every star and group, the bound the kernel forms in
    FP32 (times its 1 - 2^-16 slack) stays below the smallest accumulator of the group's models"""
    from beast_b200 import gridprep

    g, nz, obs = small
    st = FitSetup(g, nz, ["Av", "M_ini", "logT"], max_nbins=20)
    nb, nm = st.nb, st.n_models
    ng = st.ld_model // gridprep.GROUP
    assert st.boxes.shape == (ng, gridprep.box_floats(nb)) and st.boxes.dtype == np.float32
    rows = st.tiles.transpose(1, 0, 2).reshape(2 * nb + 1, -1)
    rng = np.random.default_rng(5)
    stars = np.concatenate([obs.fluxes, obs.fluxes * rng.uniform(0.2, 5.0, obs.fluxes.shape),
                            -obs.fluxes[:2], np.zeros((1, nb))])
    f32 = np.float32
    for f in stars:
        fs = (f / st.band_scale).astype(f32)
        acc = rows[2 * nb].astype(np.float64).copy()
        for b in range(nb):
            acc += (fs[b].astype(np.float64) * rows[b] + rows[nb + b]) ** 2
        one = 3 * nb + 1
        # the group's box, then the boxes of its two halves
        for which, size in ((0, gridprep.GROUP), (1, gridprep.GROUP // 2), (2, gridprep.GROUP // 2)):
            per = acc.reshape(ng, gridprep.GROUP)
            gmin = (per if which == 0 else per[:, (which - 1) * size:which * size]).min(axis=1)
            bx = st.boxes[:, which * one:(which + 1) * one]
            lb = bx[:, 3 * nb].copy()
            for b in range(nb):
                u = np.maximum(np.maximum(bx[:, b] - fs[b], fs[b] - bx[:, nb + b]), f32(0))
                d = (u * bx[:, 2 * nb + b]).astype(f32)
                lb = (d.astype(np.float64) * d + lb).astype(f32)   # fmaf
            assert np.all(lb.astype(np.float64) * 0.99998474 <= gmin * (1 + 1e-6) + 1e-30)
    # the boxes are not vacuous: a star far from a group is bounded away from it
    assert np.median(lb) > 10.0


def test_bin_constant_parameters_regroup_the_expectation(small):
    """parameters whose bins each hold one value: sum_m w_m q_m = sum_b (sum_{m in b} w_m) q_b, the
    identity behind taking their expectation from the histogram (fit.py:743 vs pdf1d.py:142-147)"""
    g, nz, obs = small
    keys = fo.default_keys(g)
    with pytest.warns(UserWarning, match="zero weight"):
        st = FitSetup(g, nz, list(keys), max_nbins=6)
    assert st.q_const.any() and not st.q_const.all()
    rng = np.random.default_rng(5)
    idx = np.sort(rng.choice(st.n_models, size=40, replace=False))
    w = rng.random(40)
    for k in range(st.n_q):
        col = st.qcols[k][st.g0_indxs]
        bn = st.bins[k].bin_num[st.g0_indxs]
        a, b = st.bin_off[k], st.bin_off[k + 1]
        if st.q_const[k]:
            # every model of a bin carries the bin's value (stored relative to the mid value) ...
            np.testing.assert_allclose(st.q_lin[k, 0] + st.bin_q[a:b][bn], col, rtol=4e-16, atol=0)
            # ... so the histogram gives the expectation
            hist = np.bincount(bn[idx], weights=w, minlength=b - a)
            got = st.q_lin[k, 0] + float(hist @ st.bin_q[a:b]) / w.sum()
            want = fo.expectation(col[idx], weights=w)
            assert abs(got - want) <= 1e-12 * max(1.0, abs(want))
        else:
            per_bin = [np.ptp(col[bn == v]) for v in np.unique(bn)]
            assert max(per_bin) > 0  # some bin mixes values: the per-model table is needed
    # a single-valued parameter comes out exactly
    single = [k for k in range(st.n_q) if st.bins[k].nbins == 1]
    for k in single:
        assert st.q_const[k] and st.bin_q[st.bin_off[k]] == 0.0


def test_error_conventions(small):
    g, nz, obs = small
    bad = dict(nz)
    bad["completeness"] = nz["completeness"].copy()
    bad["completeness"][0, 0] = 0.0
    with pytest.raises(ValueError, match="models with zero completeness present"):
        FitSetup(g, bad, ["Av"])
    with pytest.raises(KeyError, match="pdf2d_param_list cannot be None"):
        FitSetup(g, nz, ["Av"], make_pdf2d=True)
    from beast_b200 import fit

    with pytest.raises(KeyError, match='Key "nope" not recognized'):
        fit.summary_table_memory(obs, nz, g, keys=["Av", "nope"])
    with pytest.raises(KeyError, match="pdf2d_param_list cannot be None"):
        fit.summary_table_memory(obs, nz, g, keys=["Av"], pdf2d_outname="x.fits")
    cov = dict(nz)
    cov["q_norm"] = np.zeros(len(g))
    cov["icov_diag"] = nz["error"] ** -2
    cov["icov_offdiag"] = np.zeros((len(g), 15))
    assert FitSetup(g, cov, ["Av"]).full_cov  # fit.py:432-444: switches on the three arrays
    assert not FitSetup(g, cov, ["Av"], use_full_cov_matrix=False).full_cov


def test_fits_round_trip(tmp_path):
    a = np.arange(12.0).reshape(3, 4)
    b = np.arange(24.0).reshape(2, 3, 4)
    p = str(tmp_path / "img.fits")
    fitsio.write_image_file(p, [a, b], ["Av", "Av+logT"])
    hdus = fitsio.read_fits(p)
    assert hdus[0][1].shape == (2, 2) and not hdus[0][1].any()
    assert hdus[1][0]["EXTNAME"] == "Av" and hdus[1][0]["XTENSION"] == "IMAGE"
    np.testing.assert_array_equal(hdus[1][1], a)
    np.testing.assert_array_equal(hdus[2][1], b)
    assert hdus[2][0]["NAXIS1"] == 4 and hdus[2][0]["NAXIS3"] == 2
    t = str(tmp_path / "tab.fits")
    cols = {"Name": np.array(["PHAT J1", "x"]), "Av_Best": np.array([0.5, 1.5]),
            "Pmax_indx": np.array([7, 9])}
    fitsio.write_table_file(t, [cols, {"filternames": np.array(["F1", "F22"]), "wavelengths": np.array([1., 2.])}])
    hd = fitsio.read_fits(t)
    assert list(hd[1][1]) == ["Name", "Av_Best", "Pmax_indx"]
    assert list(hd[1][1]["Name"]) == ["PHAT J1", "x"] and hd[1][1]["Pmax_indx"].dtype.kind == "i"
    np.testing.assert_array_equal(hd[1][1]["Av_Best"], [0.5, 1.5])
    assert list(hd[2][1]["filternames"]) == ["F1", "F22"]
    import os

    assert os.path.getsize(t) % 2880 == 0 and os.path.getsize(p) % 2880 == 0


def test_iau_names():
    from beast_b200.fit import IAU_names_and_extra_info

    obs = synth.SyntheticObservations(np.ones((2, 6)), synth.PHAT_FILTERS, ra=[10.684708, 11.0],
                                      dec=[41.26875, -0.5])
    r = IAU_names_and_extra_info(obs, surveyname="PHAT")
    assert r["Name"][0] == "PHAT J004244.3299+411607.500"
    assert r["Name"][1].endswith("-003000.000")
    assert set(synth.PHAT_FILTERS) <= set(r) and "RA" in r and "DEC" in r
    noname = IAU_names_and_extra_info(synth.SyntheticObservations(np.ones((3, 6)), synth.PHAT_FILTERS))
    assert noname["Name"] == ["noname"] * 3


def test_golden_fixture_inputs_are_self_contained():
    g, nz, obs, args, ref = util.load_golden("tiny_thr10_pdf2d")
    assert len(g) == 1920 and obs.fluxes.shape == (16, 6) and args["threshold"] == -10.0
    assert (g["weight"] == 0).sum() > 0


def test_merging_stat_tables(tmp_path):
    # beast/tools/merge_beast_stats.py:12-71
    from beast_b200.merge import merge_stats_files

    filt = {"filternames": np.array(["F1", "F2"]), "wavelengths": np.array([1.0, 2.0])}
    names = []
    for i, n in enumerate((3, 2)):
        cols = {"Name": np.array(["s%d_%d" % (i, j) for j in range(n)]), "Pmax": np.arange(n) + 10.0 * i,
                "Pmax_indx": np.arange(n) + 100 * i}
        p = str(tmp_path / ("proj_sd0-1_sub%d_stats.fits" % i))
        fitsio.write_table_file(p, [cols, filt])
        names.append(p)
    n = merge_stats_files(names, str(tmp_path / "merged"))
    assert n == 5
    hd = fitsio.read_fits(str(tmp_path / "merged_stats.fits"))
    tab = hd[1][1]
    assert list(tab["Name"]) == ["s0_0", "s0_1", "s0_2", "s1_0", "s1_1"]
    np.testing.assert_array_equal(tab["Pmax"], [0.0, 1.0, 2.0, 10.0, 11.0])
    assert list(tab["reorder_tag"]) == ["sd0-1_sub0"] * 3 + ["sd0-1_sub1"] * 2
    assert list(hd[2][1]["filternames"]) == ["F1", "F2"]
    merge_stats_files(names, str(tmp_path / "tagged"), reorder_tag_list=["a", "b"])
    assert list(fitsio.read_fits(str(tmp_path / "tagged_stats.fits"))[1][1]["reorder_tag"]) == ["a"] * 3 + ["b"] * 2


def test_host_binding_is_a_no_op_without_a_gpu():
    """bind_host_to_gpu: without a CUDA device (or NVML) nothing is changed and None comes back"""
    import os

    from beast_b200.engine import bind_host_to_gpu

    before = os.sched_getaffinity(0)
    import torch

    if not torch.cuda.is_available():
        assert bind_host_to_gpu(0) is None
    assert os.sched_getaffinity(0) == before or torch.cuda.is_available()
