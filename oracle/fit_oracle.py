"""
TEST INFRASTRUCTURE -- CPU (NumPy, float64) restatement of the BEAST fitting hot path.

This file is the *checker*, not the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import it.  ``beast_b200`` never does.

It restates, in its own structure, what the reference computes on the path
``summary_table_memory -> Q_all_memory`` (all citations relative to
/root/reference/beast):

* setup ............ fitting/fit.py:396-573        -> ``OracleSetup``
* bin layout ....... fitting/fit.py:236-310, fitting/pdf1d.py:8-117, pdf2d.py:8-120
* likelihood ....... fitting/fit_metrics/likelihood.py:24-55, 109-167 (+58-106, 170-218)
* sparse selection . fitting/fit.py:665-690, 716-728
* expectation ...... fitting/fit_metrics/common.py:126-167
* percentile ....... fitting/fit_metrics/common.py:48-123
* gen1d / gen2d .... fitting/pdf1d.py:119-148, fitting/pdf2d.py:122-148
* symlog ........... tools/symlog.py:9-28
* subgrid merge .... tools/subgridding_tools.py:189-271, 392-527, 589-627

Parity pin: ``tests/test_oracle_vs_reference.py`` runs the unmodified reference
(``oracle/ref_shim.py``) next to this file where /root/reference exists, and
``tests/golden/*.npz`` (made by ``oracle/make_golden.py`` from the reference
itself) pin it everywhere else.  The reference's only own known-answer vectors
on this path are the symlog values (tools/tests/test_symlog.py:8-20), checked in
``tests/test_oracle_golden.py``.

Two execution modes:
  fast=True   sparse-aware gen1d/gen2d (bincount) -- same numbers, used for parity
              checks at sizes where the dense scatter would take minutes;
  fast=False  the reference's own operation sequence (per-star log(ivar),
              fancy-index copies, Python ``max``, dense scatter + per-bin sums),
              used as the timed CPU baseline so that it costs what the reference
              costs.
"""

from __future__ import annotations

import math

import numpy as np

SKIP_KEYS = "osl keep weight grid_weight prior_weight fullgrid_idx stage specgrid_indx phase eep".split()
LOG_SPACED = ("M_ini", "M_act", "radius")  # fit.py:298-301
SYMLOG_C = 1e-20  # tools/symlog.py:6


def symlog(y, C=SYMLOG_C):
    """sign(y) * log10(1 + |y/C|)   (tools/symlog.py:9-28)"""
    return np.sign(y) * np.log10(1.0 + np.abs(y / C))


def default_keys(sedgrid):
    """all grid columns minus the bookkeeping ones (fit.py:1006-1011)"""
    return [k for k in sedgrid.keys() if k not in SKIP_KEYS]


# ---------------------------------------------------------------------------
# bins
# ---------------------------------------------------------------------------
class Bins1D:
    """Bin layout of one parameter and the bin number of every grid model.

    Restates pdf1d.__init__ (pdf1d.py:32-117): linear equal-width bins when the
    parameter has more unique values than bins, otherwise one bin per unique
    value with mid-point edges and mirrored outer edges.  ``bin_num`` is
    0-based (np.digitize result minus one); the reference raises if any model
    falls outside, and so do we.
    """

    def __init__(self, gridvals, nbins, logspacing=False, minval=None, maxval=None, uniqvals=None):
        vals = np.array(gridvals, dtype=float)
        self.nbins = int(nbins)
        self.logspacing = bool(logspacing)
        if uniqvals is None:
            uniqvals = np.unique(vals)
        lo = vals.min() if minval is None else minval
        hi = vals.max() if maxval is None else maxval
        if logspacing:
            lo, hi = math.log10(lo), math.log10(hi)
            vals = np.log10(vals)
            uniqvals = np.log10(uniqvals)
        delta = (hi - lo) / (self.nbins - 1) if self.nbins > 1 else 1
        if len(uniqvals) > self.nbins:
            centres = lo + np.arange(self.nbins) * delta
            edges = lo + (np.arange(self.nbins + 1) - 0.5) * delta
        else:
            centres = np.array(uniqvals)
            edges = np.zeros(self.nbins + 1)
            if self.nbins > 1:
                edges[1:-1] = 0.5 * (centres[0:-1] + centres[1:])
                edges[0] = centres[0] - (edges[1] - centres[0])
                edges[-1] = centres[-1] + (centres[-1] - edges[-2])
            else:
                edges[0] = 0.95 * centres[0]
                edges[1] = 1.05 * centres[0]
        num = np.digitize(vals, edges)
        if np.count_nonzero((num >= 1) & (num <= self.nbins)) != len(vals):
            raise ValueError(
                "Not all the physics grid model points mapped to 1d ppdf bins - should not happen"
            )
        self.bin_num = (num - 1).astype(np.int64)
        if logspacing:
            centres = np.power(10.0, centres)
            edges = np.power(10.0, edges)
        self.bin_vals = centres
        self.bin_edges = edges
        self._lists = None

    def index_lists(self):
        """per-bin model index lists, as the reference keeps them (pdf1d.py:95-103)"""
        if self._lists is None:
            self._lists = [np.where(self.bin_num == i)[0] for i in range(self.nbins)]
        return self._lists


class Bins2D:
    """pdf2d.__init__ restated (pdf2d.py:38-120): always equal-width bins."""

    def __init__(self, v1, v2, nb1, nb2, log1=False, log2=False, min1=None, max1=None, min2=None, max2=None):
        self.nb1, self.nb2 = int(nb1), int(nb2)
        t1, t2 = np.array(v1, dtype=float), np.array(v2, dtype=float)
        lo1 = t1.min() if min1 is None else min1
        hi1 = t1.max() if max1 is None else max1
        lo2 = t2.min() if min2 is None else min2
        hi2 = t2.max() if max2 is None else max2
        if log1:
            lo1, hi1, t1 = math.log10(lo1), math.log10(hi1), np.log10(t1)
        if log2:
            lo2, hi2, t2 = math.log10(lo2), math.log10(hi2), np.log10(t2)
        d1 = (hi1 - lo1) / (self.nb1 - 1) if self.nb1 > 1 else 1
        d2 = (hi2 - lo2) / (self.nb2 - 1) if self.nb2 > 1 else 1
        c1 = lo1 + np.arange(self.nb1) * d1
        e1 = lo1 + (np.arange(self.nb1 + 1) - 0.5) * d1
        c2 = lo2 + np.arange(self.nb2) * d2
        e2 = lo2 + (np.arange(self.nb2 + 1) - 0.5) * d2
        # models outside the edges are silently dropped by the reference
        # (no bin list ever contains them, pdf2d.py:102-109); mark them -1
        n1 = np.digitize(t1, e1) - 1
        n2 = np.digitize(t2, e2) - 1
        bad = (n1 < 0) | (n1 >= self.nb1) | (n2 < 0) | (n2 >= self.nb2)
        self.bin_num = np.where(bad, -1, n1 * self.nb2 + n2).astype(np.int64)
        self.bin_vals_p1 = np.power(10.0, c1) if log1 else c1
        self.bin_vals_p2 = np.power(10.0, c2) if log2 else c2


def param_bin_spec(qname, max_nbins, sedgrid, full_model_flux, filters, grid_info_dict):
    """setup_param_bins restated (fit.py:236-310)"""
    if "_bias" in qname:
        fname = qname.replace("_wd_bias", "").replace("symlog", "")
        qvals = full_model_flux[:, filters.index(fname)]
    else:
        qvals = sedgrid[qname]
    if grid_info_dict is not None and qname in grid_info_dict:
        n_uniq = grid_info_dict[qname]["num_unique"]
        uniqvals = grid_info_dict[qname]["unique_vals"]
        lo, hi = grid_info_dict[qname]["min"], grid_info_dict[qname]["max"]
    else:
        uniqvals = np.unique(qvals)
        n_uniq = len(uniqvals)
        lo = hi = None
    nbins = max_nbins if n_uniq > max_nbins else n_uniq
    return qvals, nbins, (qname in LOG_SPACED), lo, hi, uniqvals


# ---------------------------------------------------------------------------
# statistics
# ---------------------------------------------------------------------------
def percentile(data, percentiles, weights):
    """weighted percentile on a binned pdf (fit_metrics/common.py:48-123)"""
    weights = np.asarray(weights, dtype=float)
    if np.equal(weights, 1.0).all():
        return np.percentile(data, list(percentiles))
    p = np.asarray(percentiles, dtype=float) * 0.01
    d = np.asarray(data, dtype=float)
    order = np.argsort(d)
    sd, sw = d[order], weights[order]
    aw = np.cumsum(sw)
    if not aw[-1] > 0:
        raise ValueError("Nonpositive weight sum")
    w = (aw - 0.5 * sw) / np.sum(sw)
    return np.interp(p, w, sd)


def expectation(q, weights):
    """sum(q*w)/sum(w), plain mean when every weight is exactly one (common.py:126-167)"""
    if np.equal(weights, 1.0).all():
        return np.mean(q)
    return (np.asarray(q, dtype=float) * weights).sum() / weights.sum()


# ---------------------------------------------------------------------------
# likelihood
# ---------------------------------------------------------------------------
def lnp_chi2_nm(flux, model_wbias, ivar, faithful=False, lnq=None):
    """(lnP, chi2) over all models, toothpick noise model.

    likelihood.py:109-167 with the all-False mask the fit passes (fit.py:641-645):
    chi2 = sum_b (f_b - m_b)^2 ivar_b ; lnP = -[n/2 ln 2pi - 1/2 sum_b ln ivar_b] - chi2/2.
    ``faithful`` re-creates the reference's per-star work (log of the whole ivar
    array, fancy-index copies through the mask) for timing purposes.
    """
    nb = model_wbias.shape[1]
    if faithful:
        keep = ~np.zeros(nb, dtype=bool)
        iv = ivar[:, keep]
        lnq = nb * (0.5 * np.log(2.0 * np.pi)) - 0.5 * np.sum(np.log(iv), axis=1)
        d = flux[keep] - model_wbias[:, keep]
        chi2 = np.einsum("ij,ij,ij->i", d, d, iv)
    else:
        d = flux - model_wbias
        chi2 = np.einsum("ij,ij,ij->i", d, d, ivar)
    return -lnq - 0.5 * chi2, chi2


def lnp_chi2_covar(flux, model_wbias, q_norm, icov_diag, two_icov_offdiag):
    """full-covariance likelihood (likelihood.py:58-106, 170-218)

    off-diagonals are packed row-major over the upper triangle (k<l)."""
    nb = model_wbias.shape[1]
    d = flux[None, :] - model_wbias
    chi2 = np.einsum("ij,ij,ij->i", d, d, icov_diag)
    start = 0
    for k in range(nb - 1):
        stop = start + nb - k - 1
        cross = np.einsum("ij,ij->i", two_icov_offdiag[:, start:stop], d[:, k + 1:])
        cross *= d[:, k]
        chi2 += cross
        start = stop
    lnp = -0.5 * nb * np.log(2.0 * np.pi) + q_norm - 0.5 * chi2
    return lnp, chi2


# ---------------------------------------------------------------------------
# the fit
# ---------------------------------------------------------------------------
class OracleSetup:
    """Everything Q_all_memory prepares before its star loop (fit.py:396-573)."""

    def __init__(self, sedgrid, obsmodel, qnames_in, max_nbins=200, pdf2d_param_list=None,
                 grid_info_dict=None, do_not_normalize=False, use_full_cov_matrix=True,
                 make_pdf2d=False):
        w = np.asarray(sedgrid["weight"])
        (self.g0_indxs,) = np.where(w > 0.0)
        for i in range(len(sedgrid.filters)):
            if np.any(obsmodel["completeness"][:, i] <= 0.0):
                raise ValueError("models with zero completeness present in the observation model")
        g0w = np.log(w[self.g0_indxs])
        if not do_not_normalize:
            g0w = g0w - np.log(w[self.g0_indxs].sum())
        self.g0_weights = g0w
        seds = sedgrid.seds.read() if hasattr(sedgrid.seds, "read") else sedgrid.seds
        self.full_cov = bool(
            use_full_cov_matrix and "q_norm" in obsmodel and "icov_diag" in obsmodel
            and "icov_offdiag" in obsmodel
        )
        if self.full_cov:
            self.q_norm = obsmodel["q_norm"]
            self.icov_diag = obsmodel["icov_diag"]
            self.two_icov_offdiag = 2.0 * obsmodel["icov_offdiag"]
        else:
            self.ivar = 1.0 / np.asfortranarray(obsmodel["error"]) ** 2
            nb = seds.shape[1]
            self.lnq = nb * (0.5 * np.log(2.0 * np.pi)) - 0.5 * np.sum(np.log(self.ivar), axis=1)
        self.filters = list(sedgrid.filters)
        self.qnames = list(qnames_in) + ["symlog" + f + "_wd_bias" for f in self.filters]
        self.model_wbias = np.asfortranarray(seds + obsmodel["bias"])
        self.full_model_flux = symlog(self.model_wbias)
        self.specgrid_indx = np.asarray(sedgrid["specgrid_indx"])
        self.nmodels = len(w)

        self.qvals, self.bins = [], []
        for q in self.qnames:
            vals, nbins, logsp, lo, hi, uniq = param_bin_spec(
                q, max_nbins, sedgrid, self.full_model_flux, self.filters, grid_info_dict)
            self.qvals.append(np.asarray(vals))
            self.bins.append(Bins1D(vals, nbins, logsp, lo, hi, uniq))

        self.pairs, self.bins2d = [], []
        if make_pdf2d:
            if pdf2d_param_list is None:
                raise KeyError("pdf2d_param_list cannot be None if saving 2D PDFs")
            params = [q for q in self.qnames
                      if q in pdf2d_param_list and len(np.unique(sedgrid[q])) > 1]
            for i in range(len(params)):
                for j in range(i + 1, len(params)):
                    s1 = param_bin_spec(params[i], max_nbins, sedgrid, self.full_model_flux,
                                        self.filters, grid_info_dict)
                    s2 = param_bin_spec(params[j], max_nbins, sedgrid, self.full_model_flux,
                                        self.filters, grid_info_dict)
                    self.pairs.append(params[i] + "+" + params[j])
                    self.bins2d.append(Bins2D(s1[0], s2[0], s1[1], s2[1], s1[2], s2[2],
                                              s1[3], s1[4], s2[3], s2[4]))


def fit_star(setup, flux, threshold, p=(16.0, 50.0, 84.0), fast=True):
    """One pass of the star loop body (fit.py:637-762) for one observed SED."""
    s = setup
    flux = np.asarray(flux, dtype=float)
    if s.full_cov:
        lnp, chi2 = lnp_chi2_covar(flux, s.model_wbias, s.q_norm, s.icov_diag, s.two_icov_offdiag)
    else:
        lnp, chi2 = lnp_chi2_nm(flux, s.model_wbias, s.ivar, faithful=not fast, lnq=s.lnq)
    lnp = lnp[s.g0_indxs]
    chi2 = chi2[s.g0_indxs]
    lnp += s.g0_weights
    if fast:
        top = lnp[np.isfinite(lnp)].max()
    else:
        top = max(lnp[np.isfinite(lnp)])  # the reference iterates the ndarray in Python (fit.py:670)
    (indx,) = np.where((lnp - top) > threshold)
    lnps, chi2s = lnp[indx], chi2[indx]
    log_norm = lnps.max()
    weights = np.exp(lnps - log_norm)
    weight_sum = np.sum(weights)
    weights /= weight_sum
    full = s.g0_indxs[indx]
    best_full = full[weights.argmax()]
    nq = len(s.qnames)
    out = {
        "sparse_idx": full.astype(np.int64),
        "sparse_lnp": lnps,
        "sparse_chi2": chi2s,
        "total_log_norm": log_norm + np.log(weight_sum),
        "specgrid_indx": s.specgrid_indx[best_full],
        "chi2min": chi2s.min(),
        "chi2min_indx": full[chi2s.argmin()],
        "Pmax": log_norm,
        "Pmax_indx": best_full,
        "best": np.zeros(nq),
        "exp": np.zeros(nq),
        "per": np.zeros((nq, len(p))),
        "pdf1d": [],
        "pdf2d": [],
    }
    _p = np.asarray(p, dtype=float)
    for k in range(nq):
        q = s.qvals[k]
        b = s.bins[k]
        out["best"][k] = q[best_full]
        out["exp"][k] = expectation(q[full], weights)
        if fast:
            pdf = np.bincount(b.bin_num[full], weights=weights, minlength=b.nbins)
        else:
            dense = np.zeros(s.nmodels)
            dense[full] = weights
            pdf = np.zeros(b.nbins)
            for i, lst in enumerate(b.index_lists()):
                if len(lst) > 0:
                    pdf[i] = np.sum(dense[lst])
        out["pdf1d"].append(pdf)
        if pdf.max() > 0:
            out["per"][k, :] = percentile(b.bin_vals, _p, pdf)
    for b2 in s.bins2d:
        sel = b2.bin_num[full]
        ok = sel >= 0
        pdf2 = np.bincount(sel[ok], weights=weights[ok], minlength=b2.nb1 * b2.nb2)
        out["pdf2d"].append(pdf2.reshape(b2.nb1, b2.nb2))
    return out


def fit_all(obs_fluxes, setup, threshold=-10.0, p=(16.0, 50.0, 84.0), fast=True, keep_sparse=True):
    """Run every star; return arrays laid out like the reference's save_* arguments
    (fit.py:468-478, 507-508, 569-573)."""
    fluxes = np.asarray(obs_fluxes, dtype=float)
    ns = fluxes.shape[0]
    nq = len(setup.qnames)
    res = {
        "qnames": list(setup.qnames),
        "best_vals": np.zeros((ns, nq)),
        "exp_vals": np.zeros((ns, nq)),
        "per_vals": np.zeros((ns, nq, len(p))),
        "chi2min": np.zeros(ns),
        "chi2min_indx": np.zeros(ns),
        "Pmax": np.zeros(ns),
        "Pmax_indx": np.zeros(ns),
        "specgrid_indx": np.zeros(ns),
        "total_log_norm": np.zeros(ns),
        "pdf1d": [np.zeros((ns + 1, b.nbins)) for b in setup.bins],
        "pdf2d": [np.zeros((ns + 2, b.nb1, b.nb2)) for b in setup.bins2d],
        "pdf2d_pairs": list(setup.pairs),
        "sparse": [],
    }
    for k, b in enumerate(setup.bins):
        res["pdf1d"][k][-1, :] = b.bin_vals
    for k, b in enumerate(setup.bins2d):
        res["pdf2d"][k][-2, :, :] = np.tile(b.bin_vals_p1, (b.nb2, 1)).T
        res["pdf2d"][k][-1, :, :] = np.tile(b.bin_vals_p2, (b.nb1, 1))
    for e in range(ns):
        o = fit_star(setup, fluxes[e], threshold, p, fast=fast)
        res["best_vals"][e] = o["best"]
        res["exp_vals"][e] = o["exp"]
        res["per_vals"][e] = o["per"]
        for name in ("chi2min", "chi2min_indx", "Pmax", "Pmax_indx", "specgrid_indx", "total_log_norm"):
            res[name][e] = o[name]
        for k in range(nq):
            res["pdf1d"][k][e, :] = o["pdf1d"][k]
        for k in range(len(setup.bins2d)):
            res["pdf2d"][k][e] = o["pdf2d"][k]
        if keep_sparse:
            res["sparse"].append((o["sparse_idx"], o["sparse_lnp"], o["sparse_chi2"]))
    return res


# ---------------------------------------------------------------------------
# subgrids
# ---------------------------------------------------------------------------
def uniform_slices(num_points, num_slices):
    """row slices of a grid split into subgrids (subgridding_tools.py:21-36)"""
    q, r = divmod(num_points, num_slices)
    out, start = [], 0
    for i in range(num_slices):
        stop = start + q + (1 if i < r else 0)
        out.append(slice(start, stop))
        start = stop
    return out


def subgrid_info(sedgrid, obsmodel=None):
    """min / max / unique per quantity of one subgrid (subgridding_tools.py:117-179).

    For the symlog flux quantities the reference records a stale ``unique`` from
    the previous loop; only ``num_unique > max_nbins`` matters downstream, so we
    record the flux column's own unique values (same effect: linear bins)."""
    info = {}
    for q in sedgrid.keys():
        v = np.asarray(sedgrid[q])
        info[q] = {"min": np.amin(v), "max": np.amax(v), "unique": np.unique(v)}
    if obsmodel is not None:
        fmf = symlog(sedgrid.seds + obsmodel["bias"])
        for i, f in enumerate(sedgrid.filters):
            col = fmf[:, i]
            info["symlog" + f + "_wd_bias"] = {
                "min": np.amin(col[col > -99.99]), "max": np.amax(col), "unique": np.unique(col)}
    return info


def reduce_grid_info(infos, cap_unique=1000):
    """union of per-subgrid infos (subgridding_tools.py:189-271)"""
    first = infos[0]
    out = {}
    for q in first:
        lo, hi, uniq = first[q]["min"], first[q]["max"], first[q]["unique"]
        for other in infos[1:]:
            lo = min(lo, other[q]["min"])
            hi = max(hi, other[q]["max"])
            if len(uniq) < cap_unique:
                uniq = np.union1d(uniq, other[q]["unique"])
        out[q] = {"min": lo, "max": hi, "num_unique": len(uniq), "unique_vals": uniq}
    return out


def merge_subgrid_results(results, p=(16.0, 50.0, 84.0)):
    """Arithmetic of merge_pdf1d_stats (subgridding_tools.py:392-527) on in-memory
    per-subgrid results shaped like ``fit_all`` output."""
    ng = len(results)
    ns = len(results[0]["Pmax"])
    qnames = results[0]["qnames"]
    logw = np.stack([r["total_log_norm"] for r in results], axis=1)
    maxw = logw[np.arange(ns), np.argmax(logw, axis=1)]
    w = np.exp(logw - maxw[:, None])
    merged = {"qnames": list(qnames)}
    pdfs = []
    for k in range(len(qnames)):
        acc = np.zeros_like(results[0]["pdf1d"][k])
        acc[-1, :] = results[0]["pdf1d"][k][-1, :]
        for g in range(ng):
            acc[:-1, :] += results[g]["pdf1d"][k][:ns, :] * w[:, [g]]
        norm = np.sum(acc[:-1, :], axis=1, keepdims=True)
        nz = norm[:, 0] > 0
        acc[:-1][nz, :] /= norm[nz]
        pdfs.append(acc)
    merged["pdf1d"] = pdfs
    pm = np.stack([r["Pmax"] for r in results], axis=1)
    owner = pm.argmax(axis=1)
    rows = np.arange(ns)
    merged["best_gridsub_tag"] = owner
    merged["best_vals"] = np.stack([r["best_vals"] for r in results], axis=0)[owner, rows]
    ex = np.zeros_like(results[0]["exp_vals"])
    tw = np.zeros(ns)
    for g in range(ng):
        ex += results[g]["exp_vals"] * w[:, [g]]
        tw += w[:, g]
    merged["exp_vals"] = ex / tw[:, None]
    per = np.zeros((ns, len(qnames), len(p)))
    for k in range(len(qnames)):
        centres = pdfs[k][-1]
        for e in range(ns):
            row = pdfs[k][e]
            if row.max() > 0:
                for i, pv in enumerate(p):
                    per[e, k, i] = percentile(centres, [int(pv)], row)[0]
    merged["per_vals"] = per
    merged["chi2min"] = np.stack([r["chi2min"] for r in results], axis=1).min(axis=1)
    merged["Pmax"] = pm.max(axis=1)
    merged["Pmax_indx"] = np.stack([r["Pmax_indx"] for r in results], axis=1).astype(int)[rows, owner]
    merged["total_log_norm"] = np.log(w.sum(axis=1)) + maxw
    return merged


def merge_sparse_lnp(sparse_per_subgrid, threshold=-10.0):
    """merge_lnp arithmetic (subgridding_tools.py:589-627): concatenate per star,
    keep entries within ``threshold`` of the merged maximum."""
    ng = len(sparse_per_subgrid)
    ns = len(sparse_per_subgrid[0])
    out = []
    for e in range(ns):
        idx = np.concatenate([sparse_per_subgrid[g][e][0] for g in range(ng)])
        lnp = np.concatenate([sparse_per_subgrid[g][e][1] for g in range(ng)])
        sub = np.concatenate([np.full(len(sparse_per_subgrid[g][e][0]), g) for g in range(ng)])
        if threshold is not None:
            keep = np.where((lnp - lnp.max()) > threshold)[0]
            idx, lnp, sub = idx[keep], lnp[keep], sub[keep]
        out.append((idx, lnp, sub))
    return out
