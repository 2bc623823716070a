"""
Host-side setup of a fit: everything ``Q_all_memory`` prepares before its star
loop (reference ``beast/fitting/fit.py:396-573``), laid out for the kernels.

What is computed here, once per (grid, noise model, parameter list):

* zero-weight filter ``g0_indxs`` and the log prior weights (fit.py:402, 411-415)
* completeness check (fit.py:404-409)
* ``model_seds_with_bias``, ``ivar`` and the symlog "full model flux" parameters
  (fit.py:444, 456-465; tools/symlog.py:9-28)
* the 1D bin layout of every parameter and the bin number of every model
  (fit.py:236-310, pdf1d.py:32-117) -- as one byte per (model, parameter) instead
  of the reference's per-bin index lists
* optional 2D bin layouts (fit.py:511-573, pdf2d.py:38-120)
* the FP32 tiles of the dense pass and the FP64 band-major arrays of the exact pass
  (layouts: include/beast_b200.h)

Nothing here touches CUDA; ``engine.FitEngine`` uploads the arrays.
"""

from __future__ import annotations

import math
import os
import warnings

import numpy as np

from ._cabi import CHUNK, GROUP, MAX_PARAMS, TILE_MODELS

SKIP_KEYS = "osl keep weight grid_weight prior_weight fullgrid_idx stage specgrid_indx phase eep".split()
LOG_SPACED_PARAMS = {"M_ini", "M_act", "radius"}
SYMLOG_THRESHOLD = 1e-20
FP32_U = 2.0**-24
ERR_SAFETY = 4.0
# every COARSE_STRIDE-th model makes up the coarse tiles (measured at C2: 32 -> presort 5.3 + dense 44.2 ms,
# 64 -> 2.9 + 44.7, 128 -> 1.6 + 46.3, 256 -> 1.1 + 47.2)
COARSE_STRIDE = int(os.environ.get("BF_COARSE_STRIDE", "64"))


def symlog(y, C=SYMLOG_THRESHOLD):
    """symmetric log used for the flux parameters (tools/symlog.py:9-28)"""
    return np.sign(y) * np.log10(1.0 + np.abs(y / C))


class ParamBins:
    """bin centres + 0-based bin number per model of one parameter (pdf1d.py:32-117)"""

    __slots__ = ("name", "nbins", "bin_vals", "bin_num")

    def __init__(self, name, values, nbins, logspacing, minval, maxval, uniqvals):
        v = np.array(values, dtype=np.float64)
        lo = v.min() if minval is None else minval
        hi = v.max() if maxval is None else maxval
        uniq = np.asarray(uniqvals, dtype=np.float64)
        if logspacing:
            lo, hi = math.log10(lo), math.log10(hi)
            v = np.log10(v)
            uniq = np.log10(uniq)
        step = (hi - lo) / (nbins - 1) if nbins > 1 else 1
        if len(uniq) > nbins:
            centres = lo + np.arange(nbins) * step
            edges = lo + (np.arange(nbins + 1) - 0.5) * step
        else:
            centres = np.array(uniq)
            edges = np.zeros(nbins + 1)
            if nbins > 1:
                edges[1:-1] = 0.5 * (centres[:-1] + centres[1:])
                edges[0] = centres[0] - (edges[1] - centres[0])
                edges[-1] = centres[-1] + (centres[-1] - edges[-2])
            else:
                edges[0], edges[1] = 0.95 * centres[0], 1.05 * centres[0]
        num = np.digitize(v, edges)
        if np.any((num < 1) | (num > nbins)):
            raise ValueError(
                "Not all the physics grid model points mapped to 1d ppdf bins - should not happen"
            )
        self.name = name
        self.nbins = int(nbins)
        self.bin_num = (num - 1).astype(np.int32)
        self.bin_vals = np.power(10.0, centres) if logspacing else centres


class PairBins:
    """equal-width 2D bins of a parameter pair (pdf2d.py:38-120); -1 = outside every cell"""

    __slots__ = ("name", "nb1", "nb2", "bin_vals_p1", "bin_vals_p2", "bin_num")

    def __init__(self, name, spec1, spec2):
        def axis(spec):
            vals, nbins, logsp, lo, hi, _ = spec
            v = np.array(vals, dtype=np.float64)
            lo = v.min() if lo is None else lo
            hi = v.max() if hi is None else hi
            if logsp:
                lo, hi, v = math.log10(lo), math.log10(hi), np.log10(v)
            step = (hi - lo) / (nbins - 1) if nbins > 1 else 1
            centres = lo + np.arange(nbins) * step
            edges = lo + (np.arange(nbins + 1) - 0.5) * step
            num = np.digitize(v, edges) - 1
            return (np.power(10.0, centres) if logsp else centres), num, int(nbins)

        c1, n1, self.nb1 = axis(spec1)
        c2, n2, self.nb2 = axis(spec2)
        bad = (n1 < 0) | (n1 >= self.nb1) | (n2 < 0) | (n2 >= self.nb2)
        self.name = name
        self.bin_vals_p1, self.bin_vals_p2 = c1, c2
        self.bin_num = np.where(bad, -1, n1 * self.nb2 + n2).astype(np.int32)


def _bin_spec(qname, max_nbins, sedgrid, full_model_flux, filters, grid_info_dict):
    """values, nbins, log spacing, range override, unique values (fit.py:236-310)"""
    if "_bias" in qname:
        fname = qname.replace("_wd_bias", "").replace("symlog", "")
        vals = full_model_flux[:, filters.index(fname)]
    else:
        vals = np.asarray(sedgrid[qname])
    if grid_info_dict is not None and qname in grid_info_dict:
        info = grid_info_dict[qname]
        n_uniq, uniq, lo, hi = info["num_unique"], info["unique_vals"], info["min"], info["max"]
    else:
        uniq = np.unique(vals)
        n_uniq, lo, hi = len(uniq), None, None
    nbins = max_nbins if n_uniq > max_nbins else n_uniq
    return vals, nbins, qname in LOG_SPACED_PARAMS, lo, hi, uniq


def _pad_to(n, mult):
    return ((n + mult - 1) // mult) * mult


def box_floats(nb):
    """BF_BOX_FLOATS(nb) of include/beast_b200.h"""
    return (3 * (3 * nb + 1) + 3) // 4 * 4


def _boxes_of(rows, nm, nb, size):
    """one box (3 nb + 1 floats: mlo, mhi, rlo, c2min) per run of ``size`` consecutive models"""
    nmp = rows.shape[1]
    ng = nmp // size
    r = rows[:nb].astype(np.float64)
    a = -rows[nb:2 * nb].astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        m = a / r
    valid = np.arange(nmp) < nm
    usable = valid[None, :] & np.isfinite(m)   # r = 0: the band does not constrain, rlo = 0 covers it
    big = np.finfo(np.float32).max
    mlo = np.where(usable, m, np.inf).reshape(nb, ng, size).min(axis=2)
    mhi = np.where(usable, m, -np.inf).reshape(nb, ng, size).max(axis=2)
    rlo = np.where(valid[None, :], r, np.inf).reshape(nb, ng, size).min(axis=2)
    c2min = rows[2 * nb].reshape(ng, size).min(axis=1)
    empty = ~np.isfinite(mlo) | ~np.isfinite(mhi)
    mlo = np.where(empty, -big, mlo)
    mhi = np.where(empty, big, mhi)
    rlo = np.where(np.isfinite(rlo) & ~empty, rlo, 0.0)
    lo32 = np.clip(mlo, -big, big).astype(np.float32)
    hi32 = np.clip(mhi, -big, big).astype(np.float32)
    with np.errstate(over="ignore"):
        lo32 = np.where(lo32.astype(np.float64) > mlo, np.nextafter(lo32, np.float32(-np.inf)), lo32)
        hi32 = np.where(hi32.astype(np.float64) < mhi, np.nextafter(hi32, np.float32(np.inf)), hi32)
        # one more step outwards: the kernel's r and a are the FP32 values, but a / r above is rounded
        lo32 = np.maximum(np.nextafter(lo32, np.float32(-np.inf)), -big).astype(np.float32)
        hi32 = np.minimum(np.nextafter(hi32, np.float32(np.inf)), big).astype(np.float32)
    out = np.zeros((ng, 3 * nb + 1), dtype=np.float32)
    out[:, :nb] = lo32.T
    out[:, nb:2 * nb] = hi32.T
    out[:, 2 * nb:3 * nb] = rlo.astype(np.float32).T   # r is FP32 already: exact
    out[:, 3 * nb] = c2min
    return out


def _group_boxes(rows, nm, nb):
    """Per group of GROUP models three boxes around what the dense pass reads for them (the FP32
    tile rows r_b, -a_b, c2) -- the whole group, its first and its second half --: the range of
    m_b = a_b / r_b, the smallest r_b and the smallest c2, each rounded outwards, so that
    acc >= sum_b rlo_b^2 dist(f_b, [mlo_b, mhi_b])^2 + c2min  for every model of the box
    (bf_dense_chi2_pruned).  Padding models (c2 = 3e38) stay out of the boxes."""
    ng = rows.shape[1] // GROUP
    one = 3 * nb + 1
    out = np.zeros((ng, box_floats(nb)), dtype=np.float32)
    out[:, :one] = _boxes_of(rows, nm, nb, GROUP)
    halves = _boxes_of(rows, nm, nb, GROUP // 2)
    out[:, one:2 * one] = halves[0::2]
    out[:, 2 * one:3 * one] = halves[1::2]
    return out


class FitSetup:
    """All star-independent inputs of the kernels, as NumPy arrays."""

    def __init__(self, sedgrid, obsmodel, qnames_in, max_nbins=200, pdf2d_param_list=None,
                 make_pdf2d=False, grid_info_dict=None, do_not_normalize=False,
                 use_full_cov_matrix=True):
        weight = np.asarray(sedgrid["weight"], dtype=np.float64)
        filters = list(sedgrid.filters)
        nb = len(filters)
        if nb > 12:  # (the dense sweep is instantiated per band count, 1..12)
            raise NotImplementedError("at most 12 bands are supported (got %d)" % nb)
        (g0,) = np.where(weight > 0.0)
        for i in range(nb):
            if np.any(np.asarray(obsmodel["completeness"])[:, i] <= 0.0):
                raise ValueError("models with zero completeness present in the observation model")
        lnw = np.log(weight[g0])
        if not do_not_normalize:
            lnw = lnw - np.log(weight[g0].sum())
        if len(weight) != len(g0):
            warnings.warn("some zero weight models exist")
        # the noise model carries full covariance matrices -> N_covar_logLikelihood (fit.py:432-444)
        self.full_cov = bool(use_full_cov_matrix and "q_norm" in obsmodel
                             and "icov_diag" in obsmodel and "icov_offdiag" in obsmodel)
        if self.full_cov and not (2 <= nb <= 8):
            raise NotImplementedError("the full-covariance kernels are built for 2..8 bands")

        seds = sedgrid.seds.read() if hasattr(sedgrid.seds, "read") else np.asarray(sedgrid.seds)
        err = np.asarray(obsmodel["error"], dtype=np.float64)
        msb_full = seds + np.asarray(obsmodel["bias"], dtype=np.float64)
        if self.full_cov:
            ivar_full = np.asarray(obsmodel["icov_diag"], dtype=np.float64)  # diagonal of C^-1
            two_off_full = 2.0 * np.asarray(obsmodel["icov_offdiag"], dtype=np.float64)
            q_norm = np.asarray(obsmodel["q_norm"], dtype=np.float64)
        else:
            ivar_full = 1.0 / err**2
        full_model_flux = symlog(msb_full)

        self.filters = filters
        self.lamb = np.asarray(sedgrid.lamb)
        self.nb = nb
        self.n_full = len(weight)
        self.qnames = list(qnames_in) + ["symlog" + f + "_wd_bias" for f in filters]
        nq = len(self.qnames)
        if nq > MAX_PARAMS:
            raise NotImplementedError("at most %d parameters are supported (got %d)" % (MAX_PARAMS, nq))
        self.specgrid_indx = np.asarray(sedgrid["specgrid_indx"])

        # ---- 1D bins -------------------------------------------------------------
        self.qcols, self.bins = [], []
        for q in self.qnames:
            vals, nbins, logsp, lo, hi, uniq = _bin_spec(
                q, max_nbins, sedgrid, full_model_flux, filters, grid_info_dict)
            if nbins > 256:
                raise NotImplementedError(
                    "parameter %s needs %d bins; the kernels store bin numbers in one byte "
                    "(max_nbins <= 256)" % (q, nbins))
            self.qcols.append(np.asarray(vals, dtype=np.float64))
            self.bins.append(ParamBins(q, vals, nbins, logsp, lo, hi, uniq))
        self.bin_off = np.zeros(nq + 1, dtype=np.int32)
        self.bin_off[1:] = np.cumsum([b.nbins for b in self.bins])
        self.total_bins = int(self.bin_off[-1])
        self.max_bins = int(max(b.nbins for b in self.bins))
        self.bin_val = np.concatenate([b.bin_vals for b in self.bins]).astype(np.float64)

        # ---- 2D bins -------------------------------------------------------------
        self.pairs = []
        if make_pdf2d:
            if pdf2d_param_list is None:
                raise KeyError("pdf2d_param_list cannot be None if saving 2D PDFs")
            params = [q for q in self.qnames
                      if q in pdf2d_param_list and len(np.unique(sedgrid[q])) > 1]
            specs = {q: _bin_spec(q, max_nbins, sedgrid, full_model_flux, filters, grid_info_dict)
                     for q in params}
            for i in range(len(params)):
                for j in range(i + 1, len(params)):
                    self.pairs.append(
                        PairBins(params[i] + "+" + params[j], specs[params[i]], specs[params[j]]))

        # ---- drop models the kernels cannot rank (non-finite constants) -----------
        if self.full_cov:
            # lnP = (-0.5 nb ln 2pi + q_norm) - 0.5 chi2 (likelihood.py:206-216); stored negated so
            # that the kernels use one expression for both likelihoods
            lnq = -(-0.5 * nb * np.log(2.0 * np.pi) + q_norm[g0])
        else:
            lnq = nb * (0.5 * np.log(2.0 * np.pi)) - 0.5 * np.sum(np.log(ivar_full[g0]), axis=1)
        c = -lnq + lnw
        ok = np.isfinite(c) & np.isfinite(msb_full[g0]).all(axis=1) & np.isfinite(ivar_full[g0]).all(axis=1)
        if self.full_cov:
            ok &= np.isfinite(two_off_full[g0]).all(axis=1)
        if not ok.all():
            warnings.warn("%d models with non-finite noise model / flux values are ignored"
                          % int((~ok).sum()))
            g0, lnw, lnq, c = g0[ok], lnw[ok], lnq[ok], c[ok]
        if len(g0) == 0:
            raise ValueError("no model with positive weight and finite noise model")
        self.g0_indxs = g0.astype(np.int64)
        nm = len(g0)
        nmp = _pad_to(nm, TILE_MODELS)
        self.n_models, self.ld_model = nm, nmp
        self.n_tiles = nmp // TILE_MODELS
        self.n_groups = nmp // GROUP

        # ---- exact FP64 arrays, band-major --------------------------------------
        msb = np.zeros((nb, nmp))
        ivar = np.zeros((nb, nmp))
        msb[:, :nm] = msb_full[g0].T
        ivar[:, :nm] = ivar_full[g0].T
        self.msb, self.ivar = msb, ivar
        self.lnq = np.zeros(nmp)
        self.lnq[:nm] = lnq
        self.lnw = np.zeros(nmp)
        self.lnw[:nm] = lnw

        npair = nb * (nb - 1) // 2
        if self.full_cov:
            cov_off = np.zeros((max(npair, 1), nmp))
            cov_off[:npair, :nm] = two_off_full[g0].T
            self.cov_off = cov_off
        else:
            self.cov_off = None

        # ---- FP32 tiles of the dense pass ---------------------------------------
        scale = np.median(np.abs(msb_full[g0]), axis=0)
        scale = np.where((scale > 0) & np.isfinite(scale), scale, 1.0)
        self.band_scale = scale
        self.c_off = float(c.max())
        if self.full_cov:
            nrows = 2 * nb + npair + 1
            rows = np.zeros((nrows, nmp), dtype=np.float32)
            rows[:nb, :nm] = (-msb_full[g0] / scale[None, :]).T
            rows[nb:2 * nb, :nm] = (ivar_full[g0] * scale[None, :] ** 2).T
            iu = np.triu_indices(nb, k=1)
            rows[2 * nb:2 * nb + npair, :nm] = (two_off_full[g0] * (scale[iu[0]] * scale[iu[1]])[None, :]).T
            a = msb_full[g0] * np.sqrt(np.abs(ivar_full[g0]))
        else:
            nrows = 2 * nb + 1
            inv_sigma = 1.0 / np.abs(err[g0])  # (nm, nb); error may be stored negative
            a = msb_full[g0] * inv_sigma
            rows = np.zeros((nrows, nmp), dtype=np.float32)
            rows[:nb, :nm] = (inv_sigma * scale[None, :]).T
            rows[nb:2 * nb, :nm] = (-a).T
        rows[nrows - 1, :] = 3.0e38
        rows[nrows - 1, :nm] = -2.0 * (c - self.c_off)
        self.tiles = np.ascontiguousarray(
            rows.reshape(nrows, self.n_tiles, TILE_MODELS).transpose(1, 0, 2))
        # every COARSE_STRIDE-th model, for the cheap sweep that orders the stars of a batch
        sub = rows[:, :nm:COARSE_STRIDE]
        ncp = _pad_to(sub.shape[1], TILE_MODELS)
        crow = np.zeros((nrows, ncp), dtype=np.float32)
        crow[nrows - 1, :] = 3.0e38
        crow[:, :sub.shape[1]] = sub
        self.tiles_coarse = np.ascontiguousarray(
            crow.reshape(nrows, ncp // TILE_MODELS, TILE_MODELS).transpose(1, 0, 2))
        self.boxes = None if self.full_cov else _group_boxes(rows, nm, nb)
        # the band whose flux orders the stars of a batch: the one the models are most significant
        # in (stars differ most there in units of the noise, so neighbours rule out the same groups)
        self.sort_band = int(np.argmax(np.median(np.abs(a), axis=0)))
        a_max = float(np.abs(a).max())
        # cross terms: every residual meets up to nb-1 others, with |O_kl| <= sqrt(D_k D_l)
        cov_factor = float(nb) if self.full_cov else 1.0
        self.err_a = float(ERR_SAFETY * cov_factor * 6.0 * FP32_U * math.sqrt(nb) * a_max)
        self.err_b = float(ERR_SAFETY * cov_factor * 16.0 * FP32_U)

        # ---- chunk-major parameter tables ---------------------------------------
        nch = nmp // CHUNK
        btab = np.zeros((nq, nmp), dtype=np.uint8)
        # parameter values for the expectations in 32-bit fixed point over each parameter's
        # range: q = q_lin[k,0] + q_lin[k,1] * fix, |rounding| <= range * 2^-33
        qtab = np.zeros((nq, nmp), dtype=np.int32)
        self.q_lin = np.zeros((nq, 2), dtype=np.float64)
        # parameters whose bins each hold a single value (grid axes in unique-value bins): the
        # expectation sum regroups by bin, sum_m w_m q_m = sum_b (sum_{m in b} w_m) q_b, so it is
        # taken from the histogram and the per-model table is not read for them
        self.q_const = np.zeros(nq, dtype=np.uint8)
        self.bin_q = np.zeros(self.total_bins, dtype=np.float64)
        for k in range(nq):
            btab[k, :nm] = self.bins[k].bin_num[g0]
            btab[k, nm:] = btab[k, nm - 1]  # padding rows repeat the last model (never members)
            col = self.qcols[k][g0]
            lo, hi = float(col.min()), float(col.max())
            scale = (hi - lo) / 4294967294.0
            self.q_lin[k] = (0.5 * (lo + hi), scale)
            if scale > 0:
                qtab[k, :nm] = np.rint((col - self.q_lin[k, 0]) / scale).astype(np.int64).astype(np.int32)
            rep = np.zeros(self.bins[k].nbins, dtype=np.float64)
            rep[btab[k, :nm]] = col
            if np.array_equal(rep[btab[k, :nm]], col):
                self.q_const[k] = 1
                occupied = np.zeros(self.bins[k].nbins, dtype=bool)
                occupied[btab[k, :nm]] = True
                self.bin_q[self.bin_off[k]:self.bin_off[k + 1]] = np.where(occupied, rep - self.q_lin[k, 0], 0.0)
            qtab[k, nm:] = qtab[k, nm - 1]
        self.bin_tab = np.ascontiguousarray(btab.reshape(nq, nch, CHUNK).transpose(1, 0, 2))
        self.q_tab = np.ascontiguousarray(qtab.reshape(nq, nch, CHUNK).transpose(1, 0, 2))
        self._class_split(btab, qtab, nch)
        if self.pairs:
            pb = np.full((len(self.pairs), nmp), -1, dtype=np.int32)
            for p, pair in enumerate(self.pairs):
                pb[p, :nm] = pair.bin_num[g0]
            self.pair_bin = pb
            self.pair_off = np.zeros(len(self.pairs) + 1, dtype=np.int64)
            self.pair_off[1:] = np.cumsum([pr.nb1 * pr.nb2 for pr in self.pairs])
        else:
            self.pair_bin = None
            self.pair_off = None

    def _class_split(self, btab, qtab, nch):
        """Tables of the split histogram kernel (csrc/split.cu).

        A parameter is *slow* when its value is the same for the 8 models of every chunk
        (on a BEAST grid: everything that depends on the stellar model only, since the dust
        axes run innermost), otherwise *fast*.  Slow parameters take one histogram update and
        one expectation term per chunk (the chunk's weight sum), fast ones one per model.
        The split needs at most 32 parameters per class; beyond that the kernels that treat
        all parameters alike are used."""
        nq = btab.shape[0]
        b3 = btab.reshape(nq, nch, CHUNK)
        q3 = qtab.reshape(nq, nch, CHUNK)
        slow = np.array([bool((b3[k] == b3[k][:, :1]).all() and (q3[k] == q3[k][:, :1]).all())
                         for k in range(nq)])
        self.fast_k = np.flatnonzero(~slow).astype(np.int32)
        self.slow_k = np.flatnonzero(slow).astype(np.int32)
        nf, nsl = len(self.fast_k), len(self.slow_k)
        self.split_ok = nf <= 32 and nsl <= 32
        self.fast_pad = 16 if nf <= 16 else 32
        self.slow_pad = 16 if nsl <= 16 else 32
        if not self.split_ok:
            self.fast_bin = self.fast_q = self.slow_bin = self.slow_q = self.slow_run = None
            self.fast_rng = None
            return
        fb = np.zeros((nch if nf else 1, self.fast_pad, CHUNK), dtype=np.uint8)
        fq = np.zeros((nch if nf else 1, 2, self.fast_pad, CHUNK // 2), dtype=np.int32)
        for i, k in enumerate(self.fast_k):
            fb[:, i, :] = b3[k]
            fq[:, :, i, :] = q3[k].reshape(nch, 2, CHUNK // 2)
        sb = np.zeros((nch if nsl else 1, self.slow_pad), dtype=np.uint8)
        sq = np.zeros((nch if nsl else 1, self.slow_pad), dtype=np.int32)
        for i, k in enumerate(self.slow_k):
            sb[:, i] = b3[k][:, 0]
            sq[:, i] = q3[k][:, 0]
        self.fast_bin, self.fast_q, self.slow_bin, self.slow_q = fb, fq, sb, sq
        # bin range of every chunk's 8 models, per fast parameter: lo | hi << 8
        self.fast_rng = (fb.min(axis=2).astype(np.uint16) | (fb.max(axis=2).astype(np.uint16) << 8))
        # runs of chunks with the same slow values inside a group of 8 chunks: bit c of
        # slow_run[g] = chunk c starts a run (bit 0 always).  The weights kernel sums the weights
        # per run, so the slow parameters take one update per run, not per chunk.
        start = np.ones(nch, dtype=bool)
        if nsl and nch > 1:
            start[1:] = (sb[1:] != sb[:-1]).any(axis=1) | (sq[1:] != sq[:-1]).any(axis=1)
        start[::GROUP // CHUNK] = True
        bits = (start.reshape(-1, GROUP // CHUNK) << np.arange(GROUP // CHUNK)[None, :]).sum(axis=1)
        self.slow_run = bits.astype(np.uint8)

    @property
    def n_q(self):
        return len(self.qnames)
