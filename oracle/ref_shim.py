"""
TEST INFRASTRUCTURE -- runs the *unmodified* reference fit loop in this container.

``/root/reference/beast/fitting/fit.py`` cannot be imported as shipped here:
it needs ``tables``, ``numexpr``, ``astropy`` and ``beast.physicsmodel.grid``
(h5py) at import time (fit.py:6-20).  None of them is on the in-memory path we
exercise, so placeholder modules are registered in ``sys.modules`` before the
import and the four file writers are replaced by capture functions.  Nothing in
/root/reference is edited or copied; the reference's own ``Q_all_memory``
(fit.py:313-833) executes verbatim and its outputs are returned as arrays.

This module only works where /root/reference exists (the build container).  It
is used by ``oracle/make_golden.py`` to produce ``tests/golden/*.npz`` and by the
CPU tests that pin ``oracle/fit_oracle.py`` against the reference; it is never
imported by the product (``beast_b200``), by ``-m gpu`` tests, ``smoke()`` or
``bench.py``.
"""

from __future__ import annotations

import contextlib
import io
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("BEAST_REFERENCE_ROOT", "/root/reference")
if not os.path.isfile(os.path.join(REFERENCE_ROOT, "beast", "fitting", "fit.py")):
    # the GPU box: the files oracle/make_ref.py took from the reference tree
    _vendored = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
    if os.path.isfile(os.path.join(_vendored, "beast", "fitting", "fit.py")):
        REFERENCE_ROOT = _vendored


def reference_available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "beast", "fitting", "fit.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


_fit_module = None


def load_reference_fit():
    """import beast.fitting.fit from the read-only reference tree under stubs"""
    global _fit_module
    if _fit_module is not None:
        return _fit_module
    if not reference_available():
        raise RuntimeError("reference tree not present at " + REFERENCE_ROOT)
    sys.dont_write_bytecode = True
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    def _evaluate(expr):
        # numexpr.evaluate is called once, on "g0_weights - g0_weights_sum" (fit.py:415)
        frame = sys._getframe(1)
        return eval(expr, {"__builtins__": {}}, dict(frame.f_locals))  # noqa: S307

    class _NodeError(Exception):
        pass

    if "tables" not in sys.modules:
        _stub("tables", open_file=None, exceptions=types.SimpleNamespace(NodeError=_NodeError))
    if "numexpr" not in sys.modules:
        _stub("numexpr", evaluate=_evaluate)
    if "astropy" not in sys.modules:
        ap = _stub("astropy")
        ap.units = _stub("astropy.units", degree=1.0, hourangle=1.0)
        ap.coordinates = _stub("astropy.coordinates", SkyCoord=object)
        ap.io = _stub("astropy.io")
        ap.io.fits = _stub("astropy.io.fits")
        ap.table = _stub("astropy.table", Table=dict)
    if "beast.physicsmodel.grid" not in sys.modules:
        # the real module pulls in h5py/tables; SEDGrid is only used when a file name is passed
        import beast.physicsmodel  # noqa: F401  (package __init__ is empty of heavy imports)

        _stub("beast.physicsmodel.grid", SEDGrid=object)
        sys.modules["beast.physicsmodel"].grid = sys.modules["beast.physicsmodel.grid"]

    from beast.fitting import fit  # noqa: E402

    _fit_module = fit
    return fit


def run_reference_fit(
    obs,
    sedgrid,
    obsmodel,
    qnames,
    threshold=-10.0,
    max_nbins=200,
    pdf2d_param_list=None,
    grid_info_dict=None,
    do_not_normalize=False,
    use_full_cov_matrix=True,
    p=(16.0, 50.0, 84.0),
):
    """Execute the reference ``Q_all_memory`` and return everything it would save.

    Returns a dict with the positional arguments the reference handed to
    ``save_stats`` / ``save_pdf1d`` / ``save_pdf2d`` (fit.py:40-56, 128, 158).
    """
    fit = load_reference_fit()
    captured = {}

    def cap_stats(outname, stats_dict_in, best, exp, per, chi2, chi2_indx, lnp, lnp_indx,
                  specgrid_indx, total_log_norm, qn, pp, filters, wavelengths):
        captured["stats"] = dict(
            best_vals=np.array(best), exp_vals=np.array(exp), per_vals=np.array(per),
            chi2min=np.array(chi2), chi2min_indx=np.array(chi2_indx),
            Pmax=np.array(lnp), Pmax_indx=np.array(lnp_indx),
            specgrid_indx=np.array(specgrid_indx),
            total_log_norm=np.array(total_log_norm), qnames=list(qn), p=list(pp),
        )

    def cap_pdf1d(outname, vals, qn):
        captured["pdf1d"] = [np.array(v) for v in vals]
        captured["pdf1d_qnames"] = list(qn)

    def cap_pdf2d(outname, vals, pairs):
        captured["pdf2d"] = [np.array(v) for v in vals]
        captured["pdf2d_pairs"] = list(pairs)

    saved = (fit.save_stats, fit.save_pdf1d, fit.save_pdf2d, fit.save_lnp)
    fit.save_stats, fit.save_pdf1d, fit.save_pdf2d = cap_stats, cap_pdf1d, cap_pdf2d
    fit.save_lnp = lambda *a, **k: None
    try:
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            fit.Q_all_memory(
                {},
                obs,
                sedgrid,
                obsmodel,
                list(qnames),
                p=list(p),
                max_nbins=max_nbins,
                stats_outname="capture",
                pdf1d_outname="capture",
                pdf2d_outname=("capture" if pdf2d_param_list else None),
                pdf2d_param_list=pdf2d_param_list,
                grid_info_dict=grid_info_dict,
                lnp_outname=None,
                threshold=threshold,
                use_full_cov_matrix=use_full_cov_matrix,
                do_not_normalize=do_not_normalize,
            )
    finally:
        fit.save_stats, fit.save_pdf1d, fit.save_pdf2d, fit.save_lnp = saved
    return captured


def reference_sparse_set(flux, sedgrid, obsmodel, threshold=-10.0, do_not_normalize=False):
    """The reference's sparse set for one star, using its own likelihood function.

    Three lines of the reference's arithmetic around ``N_logLikelihood_NM``
    (likelihood.py:109; fit.py:402, 411-415, 444, 463, 657-670) -- avoids having
    to stub ``tables.open_file`` to capture ``save_lnp``.
    Returns (full-grid indices int64, lnp float64, chi2 float64).
    """
    load_reference_fit()
    from beast.fitting.fit_metrics.likelihood import N_covar_logLikelihood, N_logLikelihood_NM

    (g0_indxs,) = np.where(sedgrid["weight"] > 0.0)
    g0_weights = np.log(sedgrid["weight"][g0_indxs])
    if not do_not_normalize:
        g0_weights = g0_weights - np.log(sedgrid["weight"][g0_indxs].sum())
    mod = np.asfortranarray(sedgrid.seds + obsmodel["bias"])
    if "q_norm" in obsmodel and "icov_diag" in obsmodel and "icov_offdiag" in obsmodel:
        lnp, chi2 = N_covar_logLikelihood(flux, mod, obsmodel["q_norm"], obsmodel["icov_diag"],
                                          2.0 * obsmodel["icov_offdiag"])  # fit.py:440-442, 648-655
    else:
        ivar = 1.0 / np.asfortranarray(obsmodel["error"]) ** 2
        lnp, chi2 = N_logLikelihood_NM(flux, mod, ivar, mask=np.zeros(len(flux), dtype=bool))
    lnp = lnp[g0_indxs] + g0_weights
    chi2 = chi2[g0_indxs]
    (indx,) = np.where((lnp - max(lnp[np.isfinite(lnp)])) > threshold)
    return g0_indxs[indx].astype(np.int64), lnp[indx], chi2[indx]


def run_reference_trim(sedgrid, noisemodel, obsdata, sigma_fac=3.0, trunchen=False):
    """Execute the reference ``trim_models`` (fitting/trim_grid.py:10) with its grid / file
    writers replaced by captures; returns the ``fullgrid_idx`` column it would have written."""
    load_reference_fit()
    captured = {}

    class _Grid:  # stands in for SEDGrid(lamb, seds=..., grid=..., backend="memory")
        def __init__(self, lamb, seds=None, grid=None, backend=None):
            captured["seds"] = np.array(seds)
            captured["cols"] = dict(grid)
            self.header = {}

        def write(self, name):
            captured["sed_outname"] = name

    class _File:
        def __init__(self):
            self.root = None

        def create_array(self, root, name, arr):
            captured["noise_" + name] = np.array(arr)

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    sys.modules["tables"].open_file = lambda *a, **k: _File()
    sys.modules["beast.physicsmodel.grid"].SEDGrid = _Grid
    sys.modules.pop("beast.fitting.trim_grid", None)
    from beast.fitting import trim_grid

    class _G:  # the reference reads sedgrid.seds, .lamb and .grid (a table of columns)
        seds = sedgrid.seds
        lamb = sedgrid.lamb
        grid = {k: sedgrid[k] for k in sedgrid.keys()}

    with contextlib.redirect_stdout(io.StringIO()):
        trim_grid.trim_models(_G, noisemodel, obsdata, "sed_out", "noise_out", sigma_fac=sigma_fac,
                              trunchen=trunchen)
    return captured


# ---------------------------------------------------------------------------------------
# Subgrid tools (tools/subgridding_tools.py) under in-memory stand-ins for their file I/O
# ---------------------------------------------------------------------------------------
class _Col(np.ndarray):
    pass


class MemTable:
    """the slice of astropy.table.Table the reference's merge touches: ``colnames``,
    ``t[col]`` -> array, construction from a dict (subgridding_tools.py:385-390, 530-532)"""

    def __init__(self, cols=None):
        self.cols = dict(cols or {})

    @property
    def colnames(self):
        return list(self.cols.keys())

    def __getitem__(self, name):
        return np.asarray(self.cols[name])


class _MemHDU:
    def __init__(self, name, data):
        self.name, self.data = name, data


class _MemHDUList(list):
    """fits.open(...) result: hdul[1:] -> HDUs with .name, hdul[name].data; a context manager"""

    def __getitem__(self, key):
        if isinstance(key, str):
            for h in self:
                if h.name == key:
                    return h
            raise KeyError(key)
        r = list.__getitem__(self, key)
        return _MemHDUList(r) if isinstance(key, slice) else r

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def writeto(self, name, overwrite=False):
        self.written_to = name


_subgrid_module = None


def load_reference_subgridding():
    """import beast.tools.subgridding_tools from the reference tree.  Its h5py / PyTables /
    astropy readers are replaced per call by the run_reference_* functions below."""
    global _subgrid_module
    if _subgrid_module is not None:
        return _subgrid_module
    load_reference_fit()
    import beast.observationmodel.noisemodel  # noqa: F401  (empty package __init__)
    import beast.tools  # noqa: F401

    if "beast.observationmodel.noisemodel.generic_noisemodel" not in sys.modules:
        _stub("beast.observationmodel.noisemodel.generic_noisemodel", get_noisemodelcat=None)
    if "beast.tools.read_beast_data" not in sys.modules:
        _stub("beast.tools.read_beast_data", read_lnp_data=None)
        sys.modules["beast.tools"].read_beast_data = sys.modules["beast.tools.read_beast_data"]
    from beast.tools import subgridding_tools

    _subgrid_module = subgridding_tools
    return subgridding_tools


def run_reference_grid_info(subgrids, noisemodels, cap_unique=1000):
    """the reference's ``reduce_grid_info`` (subgridding_tools.py:189-271, which calls its
    ``subgrid_info`` :117-179) on in-memory subgrids: ``SEDGrid(name)`` / ``get_noisemodelcat(name)``
    hand back ``subgrids[i]`` / ``noisemodels[i]`` for the name "sub<i>"."""
    st = load_reference_subgridding()

    class _Disk:
        def __init__(self, g):
            self._g = g
            self.seds = g.seds
            self.filters = g.filters

        def keys(self):
            return self._g.keys()

        def __getitem__(self, k):
            return self._g[k]

    saved = (st.SEDGrid, st.get_noisemodelcat)
    st.SEDGrid = lambda name, backend=None: _Disk(subgrids[int(name[3:])])
    st.get_noisemodelcat = lambda name: noisemodels[int(name[3:])]
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            names = ["sub%d" % i for i in range(len(subgrids))]
            return st.reduce_grid_info(names, names if noisemodels is not None else None,
                                       nprocs=1, cap_unique=cap_unique)
    finally:
        st.SEDGrid, st.get_noisemodelcat = saved


def run_reference_merge_pdf1d_stats(pdf1d_per_subgrid, qnames, stats_per_subgrid, filters_tab=None):
    """the reference's ``merge_pdf1d_stats`` (subgridding_tools.py:274-539) verbatim, with
    ``fits.open`` / ``Table.read`` serving the given arrays and ``save_pdf1d`` /
    ``HDUList.writeto`` captured.

    pdf1d_per_subgrid[g] : list of (nobs+1, nbins) arrays in ``qnames`` order
    stats_per_subgrid[g] : dict column -> array (the stats table of subgrid g)
    Returns (merged pdf1d arrays, merged stats dict)."""
    st = load_reference_subgridding()
    captured = {}
    ng = len(pdf1d_per_subgrid)

    def fits_open(name):
        g = int(name.split("sub")[1].split("_")[0])
        return _MemHDUList([_MemHDU("PRIMARY", None)] +
                           [_MemHDU(q, np.array(a)) for q, a in zip(qnames, pdf1d_per_subgrid[g])])

    class _Table(MemTable):
        @staticmethod
        def read(name, hdu=1):
            g = int(name.split("sub")[1].split("_")[0])
            if hdu == 1:
                return MemTable(stats_per_subgrid[g])
            if filters_tab is None:
                raise ValueError("no such HDU")
            return MemTable(filters_tab)

    def table_to_hdu(tab):
        return tab

    def cap_pdf1d(outname, vals, qn):
        captured["pdf1d"] = [np.array(v) for v in vals]
        captured["pdf1d_qnames"] = list(qn)

    fits_ns = types.SimpleNamespace(open=fits_open, HDUList=_MemHDUList, table_to_hdu=table_to_hdu)
    saved = (st.fits, st.Table, st.save_pdf1d)
    st.fits, st.Table, st.save_pdf1d = fits_ns, _Table, cap_pdf1d
    # the merged stats table is appended to an HDUList: catch it there
    orig_append = _MemHDUList.append

    def append(self, item):
        if isinstance(item, MemTable) and "stats" not in captured:
            captured["stats"] = {k: np.asarray(v) for k, v in item.cols.items()}
        orig_append(self, item)

    _MemHDUList.append = append
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            st.merge_pdf1d_stats(["sub%d_pdf1d.fits" % g for g in range(ng)],
                                 ["sub%d_stats.fits" % g for g in range(ng)], re_run=True,
                                 output_fname_base="/nonexistent/merged")
    finally:
        st.fits, st.Table, st.save_pdf1d = saved
        _MemHDUList.append = orig_append
    return captured["pdf1d"], captured["stats"]


def run_reference_merge_lnp(lnp_per_subgrid, threshold=-10.0):
    """the reference's ``merge_lnp`` (subgridding_tools.py:542-658) verbatim.
    ``lnp_per_subgrid[g]`` = list over stars of (idx, lnp) arrays.  ``read_lnp_data`` (an h5py
    reader, tools/read_beast_data.py:13-65, whose ``np.full(..., np.nan, dtype=int)`` no longer
    runs on NumPy 2) is replaced by a stand-in that pads the same way (-inf lnp; index -1), and
    ``tables.open_file`` by a capture.  Returns a list over stars of dicts idx / lnp / subgrid."""
    st = load_reference_subgridding()
    out = {}

    def read_lnp_data(fname, nstars=None, shift_lnp=True):
        g = int([i for i in fname.split("_") if "gridsub" in i][0][7:])
        stars = lnp_per_subgrid[g]
        n = max(len(s[0]) for s in stars)
        vals = np.full((n, len(stars)), -np.inf)
        indxs = np.full((n, len(stars)), -1, dtype=int)
        for k, (idx, lnp) in enumerate(stars):
            vals[:len(lnp), k] = lnp
            indxs[:len(idx), k] = idx
        return {"vals": vals, "indxs": indxs}

    class _Out:
        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

        def create_group(self, root, label, title=None):
            out[label] = {}
            return label

        def create_array(self, group, name, arr):
            out[group][name] = np.array(arr)

    saved = (st.read_beast_data.read_lnp_data, st.tables.open_file)
    st.read_beast_data.read_lnp_data = read_lnp_data
    st.tables.open_file = lambda *a, **k: _Out()
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            st.merge_lnp(["x_gridsub%d_lnp.hd5" % g for g in range(len(lnp_per_subgrid))],
                         re_run=True, output_fname_base="/nonexistent/merged", threshold=threshold)
    finally:
        st.read_beast_data.read_lnp_data, st.tables.open_file = saved
    return [out["star_%d" % i] for i in range(len(lnp_per_subgrid[0]))]


def run_reference_merge_stats(tables_by_name, filters_tab=None, reorder_tag_list=None):
    """the reference's ``merge_stats_files`` (tools/merge_beast_stats.py:12-71) on in-memory
    tables: ``Table.read`` serves ``tables_by_name[file name]``; ``Column`` / ``add_column`` /
    ``vstack`` are plain stand-ins (row-wise concatenation of equal-column tables), so what is
    pinned is the function's own logic -- tag extraction from the file names, the added
    ``reorder_tag`` column, the filter table and the return value."""
    load_reference_fit()
    captured = {}

    class _T(MemTable):
        @staticmethod
        def read(name, hdu=1):
            if hdu == 1:
                return _T(tables_by_name[name])
            if filters_tab is None:
                raise ValueError("no such HDU")
            return _T(filters_tab)

        def __len__(self):
            return len(next(iter(self.cols.values()))) if self.cols else 0

        def add_column(self, col):
            self.cols[col[1]] = np.array(col[0])

    def column(values, name=None):
        return (values, name)

    def vstack(tabs):
        names = tabs[0].colnames
        return _T({k: np.concatenate([np.asarray(t.cols[k]) for t in tabs]) for k in names})

    class _HL(_MemHDUList):
        def append(self, item):
            captured.setdefault("hdus", []).append(dict(item.cols))
            list.append(self, item)

    tmod, fmod = sys.modules["astropy.table"], sys.modules["astropy.io.fits"]
    saved_t, saved_f = dict(tmod.__dict__), dict(fmod.__dict__)
    tmod.Table, tmod.Column, tmod.vstack = _T, column, vstack
    fmod.HDUList, fmod.table_to_hdu = _HL, (lambda t: t)
    sys.modules["astropy.io"].fits = fmod
    sys.modules.pop("beast.tools.merge_beast_stats", None)
    try:
        import beast.tools  # noqa: F401
        from beast.tools import merge_beast_stats

        n = merge_beast_stats.merge_stats_files(list(tables_by_name.keys()), "/nonexistent/out",
                                                reorder_tag_list=reorder_tag_list)
    finally:  # the stubs are shared with the other run_reference_* functions
        tmod.__dict__.clear()
        tmod.__dict__.update(saved_t)
        fmod.__dict__.clear()
        fmod.__dict__.update(saved_f)
        sys.modules.pop("beast.tools.merge_beast_stats", None)
    return n, captured["hdus"]


class _TimedObs:
    """duck-typed observations whose ``enumobs`` notes the time whenever the reference's star
    loop asks for the first star of a block: the loop (fit.py:637-762) is then timed block by
    block from inside the unmodified ``Q_all_memory``, setup excluded"""

    def __init__(self, fluxes, filters, block):
        self.fluxes, self.filters, self.block, self.marks = fluxes, list(filters), block, []

    def __len__(self):
        return len(self.fluxes)

    def getFilters(self):
        return self.filters

    def enumobs(self):
        import time

        for k in range(len(self.fluxes)):
            if k % self.block == 0:
                self.marks.append(time.perf_counter())
            yield k, self.fluxes[k].copy()
        self.marks.append(time.perf_counter())


def time_reference_star_loop(fluxes, sedgrid, obsmodel, qnames, block, threshold=-10.0, max_nbins=200):
    """run the unmodified reference ``Q_all_memory`` once over ``fluxes`` and return the seconds
    its star loop spent on every block of ``block`` stars (writers replaced by no-ops)"""
    fit = load_reference_fit()
    obs = _TimedObs(fluxes, sedgrid.filters, block)
    saved = (fit.save_stats, fit.save_pdf1d, fit.save_pdf2d, fit.save_lnp)
    fit.save_stats = fit.save_pdf1d = fit.save_pdf2d = fit.save_lnp = lambda *a, **k: None
    try:
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            fit.Q_all_memory({}, obs, sedgrid, obsmodel, list(qnames), p=[16.0, 50.0, 84.0],
                             max_nbins=max_nbins, stats_outname="none", pdf1d_outname="none",
                             lnp_outname=None, threshold=threshold)
    finally:
        fit.save_stats, fit.save_pdf1d, fit.save_pdf2d, fit.save_lnp = saved
    return list(np.diff(obs.marks))
