"""CUDA path vs the reference's own outputs (golden fixtures) and vs the oracle.

Everything here goes through the C ABI (libbeast_b200.so via ctypes)."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _engine(g, nz, args):
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup

    setup = FitSetup(g, nz, list(args["keys"]), max_nbins=args["max_nbins"],
                     pdf2d_param_list=args["pdf2d_param_list"],
                     make_pdf2d=args["pdf2d_param_list"] is not None,
                     do_not_normalize=args["do_not_normalize"])
    # all parameter tables resident: BF_HIST / BF_WEIGHTS / force_legacy then choose the kernels
    return setup, FitEngine(setup, sparse_pass="both")


@pytest.mark.parametrize("case", util.GOLDEN_CASES)
def test_golden_case(case):
    g, nz, obs, args, ref = util.load_golden(case)
    setup, eng = _engine(g, nz, args)
    want2d = args["pdf2d_param_list"] is not None
    res = eng.fit(obs.fluxes, threshold=args["threshold"], want_pdf2d=want2d, sparse="all")
    assert eng.launches > 0
    util.compare_to_reference(setup, res, ref, n_pdf2d=len(setup.pairs) if want2d else 0)
    util.compare_sparse(setup, res, ref["sparse_off"], ref["sparse_idx"], ref["sparse_lnp"],
                        ref["sparse_chi2"])
    if want2d:
        assert [p.name for p in setup.pairs] == ref["pdf2d_pairs"]


def test_c1_against_oracle():
    """BASELINE.json configs[0] shape (1e5 models, 6 bands) on a star subsample, vs the oracle."""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g, nz, obs = synth.make_config("C1", nstars=48, seed=77)
    keys = fo.default_keys(g)
    ost = fo.OracleSetup(g, nz, keys, max_nbins=50)
    want = fo.fit_all(obs.fluxes, ost, threshold=-10.0, fast=True)
    args = dict(keys=keys, max_nbins=50, pdf2d_param_list=None, do_not_normalize=False)
    setup, eng = _engine(g, nz, args)
    res = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    util.compare_to_reference(setup, res, want)
    off = np.zeros(len(obs) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(s[0]) for s in want["sparse"]])
    util.compare_sparse(setup, res, off, np.concatenate([s[0] for s in want["sparse"]]),
                        np.concatenate([s[1] for s in want["sparse"]]),
                        np.concatenate([s[2] for s in want["sparse"]]))


def test_batching_is_invisible():
    """splitting the catalog into batches must not change any output bit"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    a = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    b = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all", batch=7)
    for name in ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "exp_vals", "pdf1d",
                 "per_vals", "sparse_j", "sparse_lnp"):
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name), err_msg=name)


def test_device_subsample_is_a_valid_draw():
    """lnp_npts subsample drawn on the device: min(|S|, npts) distinct members of the exact
    sparse set per star, with their exact lnP (fit.py:693-709)"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    full = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    npts = 20
    sub = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=npts, lnp_rng="device", seed=3)
    again = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=npts, lnp_rng="device", seed=3)
    np.testing.assert_array_equal(sub.sparse_j, again.sparse_j)
    for e in range(len(obs)):
        fa, fb = full.sparse_off[e], full.sparse_off[e + 1]
        sa, sb = sub.sparse_off[e], sub.sparse_off[e + 1]
        assert sb - sa == min(fb - fa, npts)
        picked = sub.sparse_j[sa:sb]
        assert len(set(picked.tolist())) == sb - sa
        lookup = dict(zip(full.sparse_j[fa:fb].tolist(), full.sparse_lnp[fa:fb].tolist()))
        for j, v in zip(picked.tolist(), sub.sparse_lnp[sa:sb].tolist()):
            assert lookup[j] == v


def test_numpy_subsample_follows_reference_stream():
    """lnp_rng="numpy": the ranks are np.random.choice(|S|, npts, replace=False) in catalog
    order, the call the reference makes (fit.py:696)"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    full = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    npts = 20
    np.random.seed(1234)
    sub = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=npts, lnp_rng="numpy")
    np.random.seed(1234)
    for e in range(len(obs)):
        fa, fb = full.sparse_off[e], full.sparse_off[e + 1]
        indx = np.arange(fb - fa)
        rindx = np.random.choice(indx, size=npts, replace=False) if npts < len(indx) else indx
        sa, sb = sub.sparse_off[e], sub.sparse_off[e + 1]
        np.testing.assert_array_equal(sub.sparse_j[sa:sb], full.sparse_j[fa:fb][rindx])


@pytest.mark.parametrize("layout", ["full", "window"])
def test_histogram_layouts_agree_with_the_oracle(layout, monkeypatch):
    """200-bin tables: the full-width histogram kernel and the windowed kernel both give the
    oracle's pdfs, percentiles and expectations (BF_HIST is read at every call of
    bf_exact_accumulate; the golden cases run whichever kernel the bin tables select)."""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    monkeypatch.setenv("BF_HIST", layout)
    g, nz, obs = synth.make_config("C1", nstars=40, seed=5)
    keys = fo.default_keys(g)
    ost = fo.OracleSetup(g, nz, keys, max_nbins=200)
    want = fo.fit_all(obs.fluxes, ost, threshold=-10.0, fast=True)
    args = dict(keys=keys, max_nbins=200, pdf2d_param_list=None, do_not_normalize=False)
    setup, eng = _engine(g, nz, args)
    assert setup.q_const.any() and not setup.q_const.all()  # both expectation paths are used
    res = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    util.compare_to_reference(setup, res, want)


def test_full_width_kernel_at_100_bins(monkeypatch):
    """100-bin tables leave the full-width histogram kernel 8 warps: it then adds four models at a
    time (as the windowed kernel does); BF_HIST_SERIAL selects the one-member-at-a-time loop"""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g, nz, obs = synth.make_config("C1", nstars=24, seed=9)
    keys = fo.default_keys(g)
    want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=100), threshold=-10.0, fast=True)
    args = dict(keys=keys, max_nbins=100, pdf2d_param_list=None, do_not_normalize=False)
    monkeypatch.setenv("BF_HIST", "full")
    setup, eng = _engine(g, nz, args)
    res = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    util.compare_to_reference(setup, res, want)
    monkeypatch.setenv("BF_HIST_SERIAL", "1")
    res2 = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    util.compare_to_reference(setup, res2, want)
    np.testing.assert_allclose(res.pdf1d, res2.pdf1d, rtol=1e-12, atol=1e-300)


def test_weights_kernel_instantiations_are_bit_identical(monkeypatch):
    """the per-band-count weights kernel (all loads of a model in flight at once) performs the
    same operations in the same order as the runtime-band-count one: every output bit agrees"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    monkeypatch.setenv("BF_WEIGHTS", "star")
    a = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    monkeypatch.setenv("BF_WEIGHTS_GENERIC", "1")
    b = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    for name in ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse", "pdf1d",
                 "exp_vals", "per_vals", "sparse_j", "sparse_lnp", "sparse_chi2"):
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name), err_msg=name)


def test_pdf2d_entry_points_agree():
    """bf_pdf2d_from_weights (weights left by the sparse pass, used inside bf_fit_batch) and the
    stand-alone bf_pdf2d_accumulate (models evaluated again) build the same 2D PDFs; both match
    the reference's (golden case)"""
    import ctypes as C

    import torch

    from beast_b200 import _cabi

    g, nz, obs, args, ref = util.load_golden("tiny_thr10_pdf2d")
    setup, eng = _engine(g, nz, args)
    eng.presort = False  # the lists bf_fit_batch leaves behind are then in catalog order
    a = eng.fit(obs.fluxes, threshold=args["threshold"], want_pdf2d=True, sparse="all")
    util.compare_to_reference(setup, a, ref, n_pdf2d=len(setup.pairs))
    plan, state = eng.last_state
    n = len(obs)
    dev = eng.device
    pmax = torch.from_numpy(a.Pmax.copy()).to(dev)
    tln = torch.from_numpy(a.total_log_norm.copy()).to(dev)
    pdf2d = torch.zeros((n, int(setup.pair_off[-1])), dtype=torch.float64, device=dev)
    P = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    _cabi.call("bf_pdf2d_accumulate", C.byref(eng.grid), len(setup.pairs), P(eng.d_pair_bin),
               P(eng.d_pair_off), n, C.c_void_p(state.flux64), P(pmax), P(tln), float(args["threshold"]),
               C.c_void_p(state.offsets), C.c_void_p(state.list), P(pdf2d),
               C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    torch.cuda.synchronize(dev)
    np.testing.assert_allclose(a.pdf2d[:n], pdf2d.cpu().numpy(), rtol=1e-12, atol=1e-300)


def test_fit_batch_is_one_c_call_and_presort_is_invisible():
    """a whole fit through bf_fit_batch: with and without the presort (stars fitted in another
    order and put back) every output agrees bit for bit, sparse records included"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    a = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=20, lnp_rng="device", seed=5)
    eng.presort = False
    b = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=20, lnp_rng="device", seed=5)
    _same_everything(a, b, ALL_OUT + ("sparse_off", "n_groups_selected"))
    util.compare_to_reference(setup, a, ref)


def test_pruned_dense_sweep_leaves_the_fit_unchanged():
    """the dense sweep that skips groups no star of a warp can select: the same lists, hence the
    same fit bit for bit, on C1-sized stars (where most of the grid is skipped) and on the golden"""
    from beast_b200 import synth

    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    assert eng.prune and eng.d_boxes is not None
    a = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=20, lnp_rng="device", seed=5)
    eng.prune = False
    b = eng.fit(obs.fluxes, threshold=args["threshold"], sparse=20, lnp_rng="device", seed=5)
    _same_everything(a, b, ALL_OUT + ("sparse_off", "n_groups_selected"))

    g, nz, obs = synth.make_config("C1", nstars=3000)
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup
    from oracle import fit_oracle as fo

    setup = FitSetup(g, nz, list(fo.default_keys(g)), max_nbins=50)
    eng = FitEngine(setup)
    a = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    eng.prune = False
    b = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    _same_everything(a, b, ALL_OUT + ("sparse_off", "n_groups_selected"))


def test_pruned_dense_entry_point():
    """bf_dense_chi2_pruned against bf_dense_chi2 on the same tiles: a group is either reported as
    before or skipped (+inf), skipped only where its minimum is at or above the star's ceiling,
    and the star minima (value and group) are the same"""
    import torch

    from beast_b200 import _cabi, synth
    from beast_b200.gridprep import FitSetup

    g, nz, obs = synth.make_config("C1", nstars=2048)
    setup = FitSetup(g, nz, ["Av", "M_ini", "logT"], max_nbins=20)
    dev = torch.device("cuda")
    nb, n, ld = setup.nb, len(obs), 2048
    # stars ordered by brightness so that a warp's 256 stars are alike
    order = np.argsort(obs.fluxes.sum(axis=1))
    flux = torch.from_numpy(np.ascontiguousarray((obs.fluxes[order] / setup.band_scale).T.astype(np.float32))).to(dev)
    tiles = torch.from_numpy(setup.tiles).to(dev)
    boxes = torch.from_numpy(setup.boxes).to(dev)
    ng = setup.n_groups

    count = torch.zeros(1, dtype=torch.int64, device=dev)

    def sweep(cut):
        count.zero_()
        gmin = torch.empty((ng, ld), dtype=torch.float32, device=dev)
        smin = torch.full((ld,), 0x7f800000ffffffff, dtype=torch.int64, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        if cut is None:
            _cabi.call("bf_dense_chi2", nb, setup.n_tiles, tiles.data_ptr(), n, ld, flux.data_ptr(),
                       gmin.data_ptr(), smin.data_ptr(), 4, st)
        else:
            _cabi.call("bf_dense_chi2_pruned", nb, setup.n_tiles, tiles.data_ptr(), boxes.data_ptr(), n, ld,
                       flux.data_ptr(), cut.data_ptr(), gmin.data_ptr(), smin.data_ptr(), 4, count.data_ptr(), st)
        torch.cuda.synchronize()
        return gmin.cpu().numpy(), smin.cpu().numpy()

    full, smin_full = sweep(None)
    star_min = full.min(axis=0)
    cut = torch.from_numpy((star_min + 25.0).astype(np.float32)).to(dev)
    pruned, smin_pruned = sweep(cut)
    np.testing.assert_array_equal(smin_pruned, smin_full)
    skipped = np.isinf(pruned) & ~np.isinf(full)
    assert skipped.mean() > 0.3            # the sweep does leave things out
    # reported as before -- or, with one half of the group ruled out for the star's pair, the other
    # half's minimum, then at or above the ceiling like everything in the half left out
    cutb = np.broadcast_to(cut.cpu().numpy()[None, :], full.shape)
    differ = ~skipped & (pruned != full)
    assert np.all(pruned[differ] >= cutb[differ]) and np.all(full[differ] >= cutb[differ])
    below = full < cutb
    np.testing.assert_array_equal(pruned[below], full[below])
    # the counter: halves of groups for 32 + 32 stars at a time, so it lies between what is
    # below the ceilings and everything
    assert below.sum() * 32 <= int(count.item()) < 0.7 * ng * 64 * ld
    assert np.all(full[skipped] >= cutb[skipped])
    # a ceiling of +inf keeps everything, -1 (padding stars) skips everything
    same, _ = sweep(torch.full((ld,), float("inf"), device=dev))
    np.testing.assert_array_equal(same, full)
    assert int(count.item()) == ng * 64 * ld
    none, _ = sweep(torch.full((ld,), -1.0, device=dev))
    assert np.all(np.isinf(none))


def _same_everything(a, b, names):
    for name in names:
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name), err_msg=name)


ALL_OUT = ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse", "pdf1d", "exp_vals",
           "per_vals", "sparse_j", "sparse_lnp", "sparse_chi2")


def test_group_ordered_weights_match_star_ordered(monkeypatch):
    """the exact weights evaluated group by group (rows in registers, stars of the group in turn)
    and star by star are the same operations on the same numbers: memberships, weights, chi2
    minima agree bit for bit; the group's weight sum is added up in a different order"""
    g, nz, obs, args, ref = util.load_golden("small_thr10")
    setup, eng = _engine(g, nz, args)
    assert eng.split
    monkeypatch.setenv("BF_HIST", "window")
    a = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    monkeypatch.setenv("BF_WEIGHTS", "star")
    b = eng.fit(obs.fluxes, threshold=args["threshold"], sparse="all")
    _same_everything(a, b, ("Pmax", "Pmax_j", "chi2min", "chi2min_j", "n_sparse", "sparse_j",
                            "sparse_lnp", "sparse_chi2"))
    np.testing.assert_allclose(a.total_log_norm, b.total_log_norm, rtol=1e-14, atol=0)
    np.testing.assert_allclose(a.pdf1d, b.pdf1d, rtol=1e-13, atol=1e-300)
    # (the fast class's expectation sums: FP64 tensor-core products per group vs a serial loop)
    span = np.array([c[setup.g0_indxs].max() - c[setup.g0_indxs].min() for c in setup.qcols])
    assert np.all(np.abs(a.exp_vals - b.exp_vals) <= 1e-12 * np.abs(b.exp_vals) + 1e-12 * span)


@pytest.mark.parametrize("nbins", [50, 200])
def test_split_histogram_kernel_against_oracle_and_legacy(nbins):
    """the class-split histogram kernel (slow parameters once per chunk, fast ones two chunks per
    instruction) against the oracle, and against the kernels that treat all parameters alike"""
    from beast_b200 import synth
    from oracle import fit_oracle as fo

    g, nz, obs = synth.make_config("C1", nstars=40, seed=5)
    keys = fo.default_keys(g)
    want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=nbins), threshold=-10.0, fast=True)
    args = dict(keys=keys, max_nbins=nbins, pdf2d_param_list=None, do_not_normalize=False)
    setup, eng = _engine(g, nz, args)
    assert eng.split and len(setup.fast_k) > 0 and len(setup.slow_k) > 0
    res = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    util.compare_to_reference(setup, res, want)
    eng.force_legacy = True
    old = eng.fit(obs.fluxes, threshold=-10.0, sparse="all")
    _same_everything(res, old, ("Pmax", "Pmax_j", "chi2min", "chi2min_j", "n_sparse", "sparse_j",
                                "sparse_lnp"))
    np.testing.assert_allclose(res.total_log_norm, old.total_log_norm, rtol=1e-14, atol=0)
    np.testing.assert_allclose(res.pdf1d, old.pdf1d, rtol=1e-11, atol=1e-300)
    span = np.array([c[setup.g0_indxs].max() - c[setup.g0_indxs].min() for c in setup.qcols])
    assert np.all(np.abs(res.exp_vals - old.exp_vals) <= 1e-12 * np.abs(old.exp_vals) + 1e-9 * span)


def test_split_kernel_with_one_class_only():
    """parameter lists that are all fast (dust axes) or all slow (stellar axes)"""
    from oracle import fit_oracle as fo

    g, nz, obs, args, ref = util.load_golden("small_thr10")
    for keys in (["Av", "Rv", "f_A"], ["M_ini", "logA", "Z", "logT"]):
        a = dict(args, keys=keys)
        want = fo.fit_all(obs.fluxes, fo.OracleSetup(g, nz, keys, max_nbins=a["max_nbins"]),
                          threshold=a["threshold"], fast=True)
        setup, eng = _engine(g, nz, a)
        assert eng.split
        res = eng.fit(obs.fluxes, threshold=a["threshold"], sparse="all")
        util.compare_to_reference(setup, res, want)
