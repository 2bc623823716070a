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
