"""
Subgrid mode: the model axis is split into row slices, every slice is fitted against all
stars with ``do_not_normalize=True`` and a shared ``grid_info_dict``, and the per-star
results are merged (reference: ``beast/tools/subgridding_tools.py`` and
``docs/subgrid_parallelism.rst``).

The reference merges through files (``merge_pdf1d_stats``, subgridding_tools.py:274-539).
Here one rank holds one subgrid on its GPU and the merge is a handful of collectives on
device tensors (NCCL over NVLink) plus three small kernels (``csrc/merge.cu``):

    all_gather  total_log_norm, Pmax        -> weights exp(tln_g - max_g)      :394-406
    all_reduce  SUM  weight_g * pdf1d_g     -> normalise rows, percentiles     :421-434, 472-491
    all_reduce  SUM  weight_g * Exp_g       -> / sum_g weight_g                :462-470
    all_reduce  SUM  owner-masked Best / Pmax_indx  (owner = subgrid with the highest Pmax)
                                                                                :443-460, 506-511
    all_reduce  MIN  chi2min ; Pmax = max_g ; total_log_norm = ln sum_g w + max  :493-514

``chi2min_indx`` and ``specgrid_indx`` are dropped and ``best_gridsub_tag`` is added, as the
reference does (:523-527).  The same code merges subgrids fitted one after another on a
single GPU (``group=None`` with a list of results).
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _cabi
from .gridprep import symlog

MERGE_FIELDS = ("total_log_norm", "Pmax", "Pmax_indx", "chi2min", "best_vals", "exp_vals", "pdf1d")


def uniform_slices(num_points, num_slices):
    """row slices of the subgrids (subgridding_tools.py:21-36)"""
    q, r = divmod(num_points, num_slices)
    out, start = [], 0
    for i in range(num_slices):
        stop = start + q + (1 if i < r else 0)
        out.append(slice(start, stop))
        start = stop
    return out


def subgrid_info(sedgrid, obsmodel=None):
    """min / max / unique values of every quantity of one subgrid (subgridding_tools.py:117-179)"""
    info = {}
    for q in sedgrid.keys():
        v = np.asarray(sedgrid[q])
        info[q] = {"min": np.amin(v), "max": np.amax(v), "unique": np.unique(v)}
    if obsmodel is not None:
        seds = sedgrid.seds.read() if hasattr(sedgrid.seds, "read") else np.asarray(sedgrid.seds)
        flux = symlog(seds + np.asarray(obsmodel["bias"]))
        for i, f in enumerate(sedgrid.filters):
            col = flux[:, i]
            info["symlog" + f + "_wd_bias"] = {"min": np.amin(col[col > -99.99]), "max": np.amax(col),
                                               "unique": np.unique(col)}
    return info


def reduce_grid_info(infos, cap_unique=1000):
    """union over subgrids so that every subgrid bins identically (subgridding_tools.py:189-271)"""
    out = {}
    for q in infos[0]:
        lo, hi, uniq = infos[0][q]["min"], infos[0][q]["max"], infos[0][q]["unique"]
        for other in infos[1:]:
            lo, hi = min(lo, other[q]["min"]), max(hi, other[q]["max"])
            if len(uniq) < cap_unique:
                uniq = np.union1d(uniq, other[q]["unique"])
        out[q] = {"min": lo, "max": hi, "num_unique": len(uniq), "unique_vals": uniq}
    return out


class CudaMergeMath:
    """the arithmetic of the merge on CUDA tensors, through the C ABI (csrc/merge.cu)"""

    def __init__(self, setup, percentiles=(16.0, 50.0, 84.0)):
        import torch

        self.torch = torch
        self.setup = setup
        self.set_percentiles(percentiles)
        self._tables = None

    def set_percentiles(self, percentiles):
        """the three percentiles ``finalize`` computes (in percent)"""
        self.pct = (C.c_double * 3)(*[float(x) for x in percentiles])

    def _stream(self, t):
        return C.c_void_p(self.torch.cuda.current_stream(t.device).cuda_stream)

    def _params(self, device):
        if self._tables is None:
            s = self.setup
            self._off = self.torch.from_numpy(s.bin_off).to(device)
            self._val = self.torch.from_numpy(s.bin_val).to(device)
            self._tables = _cabi.ParamTables(s.n_q, s.total_bins, s.max_bins, 0, 0, 0,
                                             self._off.data_ptr(), self._val.data_ptr())
        return self._tables

    def weights(self, tln_all):
        """[G, ns] total_log_norm -> ([G, ns] weights, [ns] merged total_log_norm)"""
        torch = self.torch
        if not tln_all.is_cuda:
            raise RuntimeError("beast_b200 merge kernels need CUDA tensors; there is no CPU fallback")
        tln_all = tln_all.contiguous()
        g, ns = tln_all.shape
        w = torch.empty_like(tln_all)
        merged = torch.empty(ns, dtype=torch.float64, device=tln_all.device)
        _cabi.call("bf_merge_weights", g, ns, C.c_void_p(tln_all.data_ptr()), C.c_void_p(w.data_ptr()),
                   C.c_void_p(merged.data_ptr()), self._stream(tln_all))
        return w, merged

    def scale_rows(self, rows, w):
        rows = rows.contiguous().clone()
        ns = rows.shape[0]
        _cabi.call("bf_scale_rows", ns, rows.numel() // ns, C.c_void_p(w.contiguous().data_ptr()),
                   C.c_void_p(rows.data_ptr()), self._stream(rows))
        return rows

    def axpy_rows(self, w, src, dst, accumulate):
        """dst = (dst if accumulate else 0) + w[:, None] * src, in place on the device"""
        ns = src.shape[0]
        _cabi.call("bf_axpy_rows", ns, src.numel() // ns, C.c_void_p(w.contiguous().data_ptr()),
                   C.c_void_p(src.contiguous().data_ptr()), C.c_void_p(dst.data_ptr()),
                   int(bool(accumulate)), self._stream(src))

    def finalize(self, pdf):
        """normalise every parameter's pdf row and recompute the percentiles; returns per_vals"""
        torch = self.torch
        ns = pdf.shape[0]
        per = torch.empty((ns, self.setup.n_q, 3), dtype=torch.float64, device=pdf.device)
        tables = self._params(pdf.device)
        _cabi.call("bf_merge_finalize", C.byref(tables), ns, C.c_void_p(pdf.data_ptr()), self.pct,
                   C.c_void_p(per.data_ptr()), self._stream(pdf))
        return per


def merge_subgrids(local, math, group=None, rank=0, world=1):
    """Merge per-subgrid results.

    ``local``: with a process group, this rank's dict of torch tensors (MERGE_FIELDS, one row
    per star); without one, a list of such dicts (all subgrids in this process).
    ``math``: object with ``weights``, ``scale_rows``, ``finalize`` (``CudaMergeMath``).
    Returns the merged dict (+ ``best_gridsub_tag``, ``per_vals``)."""
    import torch

    if group is None and isinstance(local, (list, tuple)):
        parts = list(local)
        world = len(parts)

        def gather(name):
            return torch.stack([p[name] for p in parts], dim=0)

        def reduce_sum(fn):
            acc = fn(0)
            for g in range(1, world):
                acc = acc + fn(g)
            return acc

        def reduce_min(name):
            return torch.stack([p[name] for p in parts], 0).min(0).values

        ranks = range(world)
    else:
        import torch.distributed as dist

        parts = {rank: local}

        def gather(name):
            buf = [torch.empty_like(local[name]) for _ in range(world)]
            dist.all_gather(buf, local[name].contiguous(), group=group)
            return torch.stack(buf, dim=0)

        def reduce_sum(fn):
            t = fn(rank).contiguous()
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
            return t

        def reduce_min(name):
            t = local[name].clone()
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            return t

        ranks = [rank]

    tln_all = gather("total_log_norm")
    pmax_all = gather("Pmax")
    w_all, merged_tln = math.weights(tln_all)
    wsum = w_all.sum(0)
    pdf = reduce_sum(lambda g: math.scale_rows(parts[g]["pdf1d"], w_all[g]))
    exp = reduce_sum(lambda g: math.scale_rows(parts[g]["exp_vals"], w_all[g])) / wsum[:, None]
    per = math.finalize(pdf)
    # owner = first subgrid holding the highest Pmax (np.argmax, subgridding_tools.py:447)
    top = pmax_all.max(0).values
    idx = torch.arange(world, device=pmax_all.device)[:, None].expand_as(pmax_all)
    owner = torch.where(pmax_all == top[None, :], idx, torch.full_like(idx, world)).min(0).values
    best = reduce_sum(lambda g: torch.where((owner == g)[:, None], parts[g]["best_vals"],
                                            torch.zeros_like(parts[g]["best_vals"])))
    pidx = reduce_sum(lambda g: torch.where(owner == g, parts[g]["Pmax_indx"],
                                            torch.zeros_like(parts[g]["Pmax_indx"])))
    return {
        "pdf1d": pdf, "per_vals": per, "exp_vals": exp, "best_vals": best, "Pmax": top,
        "Pmax_indx": pidx, "best_gridsub_tag": owner, "chi2min": reduce_min("chi2min"),
        "total_log_norm": merged_tln,
    }


class SubgridFit:
    """Subgrid mode without files (docs/subgrid_parallelism.rst:77-143, on device): every process
    holds one or more row slices of the grid resident on its GPU (``engines``: FitEngine objects
    built with ``do_not_normalize=True`` and the shared ``grid_info_dict``), all processes fit the
    same batch of stars, and the per-star partial results are merged while they are still in HBM:

        all_gather   total_log_norm, Pmax, chi2min of every subgrid  -> weights, owner, minima
        local sum    weight_g * pdf1d_g, weight_g * Exp_g, owner-masked Best / Pmax_indx
                     over the subgrids of this process (bf_axpy_rows)
        reduce_scatter(SUM) by star of those sums (NCCL over NVLink): every process ends up with
                     the merged rows of its share of the batch's stars
        normalise + percentiles on that share (bf_merge_finalize)

    -- the arithmetic of merge_pdf1d_stats (subgridding_tools.py:392-527) with the reference's
    files replaced by one collective per batch.  Subgrid g of the run is local subgrid
    g - rank * len(engines) of process rank."""

    def __init__(self, engines, math, group=None, rank=0, world=1):
        self.engines, self.math = list(engines), math
        self.group, self.rank, self.world = group, rank, world
        self.collective_bytes = 0   # payload this process put into reduce_scatter / all_gather
        self.collective_ms = 0.0    # device time of the collectives when ``time_collectives``
        self.time_collectives = False
        self._cache = {}

    def _tables(self, setup, dev):
        """device copies of a subgrid's g0_indxs and parameter columns (Best values), made once"""
        import torch

        key = id(setup)
        if key not in self._cache:
            self._cache[key] = (torch.from_numpy(setup.g0_indxs).to(dev),
                                torch.from_numpy(np.stack([np.asarray(c, dtype=np.float64)
                                                           for c in setup.qcols])).to(dev))
        return self._cache[key]

    def share(self, n):
        """rows [lo, hi) of a batch of n stars that this process ends up with"""
        per = -(-n // self.world)
        return min(n, self.rank * per), min(n, (self.rank + 1) * per)

    def fit_batch(self, flux, threshold=-10.0, **kw):
        outs = [eng.fit_batch(flux, threshold, **kw) for eng in self.engines]
        return self.merge_batch(outs, [eng.setup for eng in self.engines])

    def merge_batch(self, outs, setups):
        """outs[l]: the dict of device tensors FitEngine.fit_batch returned for local subgrid l"""
        import torch
        import torch.distributed as dist

        L, W = len(outs), self.world
        n = outs[0]["Pmax"].shape[0]
        dev = outs[0]["Pmax"].device
        per = -(-n // W)
        npad = per * W
        loc = torch.stack([torch.stack([o["total_log_norm"], o["Pmax"], o["chi2min"]]) for o in outs])  # [L,3,n]
        if W > 1:
            allv = torch.empty((W,) + tuple(loc.shape), dtype=loc.dtype, device=dev)
            self._timed(lambda: dist.all_gather_into_tensor(allv, loc.contiguous(), group=self.group)
                        if dist.get_backend(self.group) != "gloo" else
                        dist.all_gather(list(allv.unbind(0)), loc.contiguous(), group=self.group))
            self.collective_bytes += loc.numel() * 8
            allv = allv.reshape(W * L, 3, n)
        else:
            allv = loc
        tln_all, pmax_all, chi2_all = allv[:, 0].contiguous(), allv[:, 1].contiguous(), allv[:, 2]
        w_all, merged_tln = self.math.weights(tln_all)
        wsum = w_all.sum(0)
        top = pmax_all.max(0).values
        G = W * L
        idx = torch.arange(G, device=dev)[:, None].expand_as(pmax_all)
        owner = torch.where(pmax_all == top[None, :], idx, torch.full_like(idx, G)).min(0).values
        nq = outs[0]["exp_vals"].shape[1]
        tb = outs[0]["pdf1d"].shape[1]
        acc_pdf = torch.zeros((npad, tb), dtype=torch.float64, device=dev)
        acc_small = torch.zeros((npad, 2 * nq + 1), dtype=torch.float64, device=dev)
        ex = torch.zeros((n, nq), dtype=torch.float64, device=dev)
        for l, (o, st) in enumerate(zip(outs, setups)):
            g = self.rank * L + l
            w = w_all[g].contiguous()
            mine = (owner == g).to(torch.float64)
            self.math.axpy_rows(w, o["pdf1d"], acc_pdf[:n], l > 0)
            self.math.axpy_rows(w, o["exp_vals"], ex, l > 0)
            g0_d, q_d = self._tables(st, dev)
            best_full = g0_d[o["Pmax_j"].clamp(min=0)]  # row of this subgrid (fit.py:719)
            best_vals = q_d[:, best_full].t()
            acc_small[:n, nq:2 * nq] += best_vals * mine[:, None]
            acc_small[:n, 2 * nq] += best_full.to(torch.float64) * mine
        acc_small[:n, :nq] = ex
        if W > 1:
            pdf = torch.empty((per, tb), dtype=torch.float64, device=dev)
            small = torch.empty((per, 2 * nq + 1), dtype=torch.float64, device=dev)
            if dist.get_backend(self.group) == "gloo":  # (CPU tests: gloo has no reduce_scatter)
                dist.all_reduce(acc_pdf, group=self.group)
                dist.all_reduce(acc_small, group=self.group)
                pdf.copy_(acc_pdf[self.rank * per:(self.rank + 1) * per])
                small.copy_(acc_small[self.rank * per:(self.rank + 1) * per])
            else:
                self._timed(lambda: (dist.reduce_scatter_tensor(pdf, acc_pdf, group=self.group),
                                     dist.reduce_scatter_tensor(small, acc_small, group=self.group)))
            self.collective_bytes += (acc_pdf.numel() + acc_small.numel()) * 8
        else:
            pdf, small = acc_pdf, acc_small
        lo, hi = self.share(n)
        m = hi - lo
        pdf, small = pdf[:m], small[:m]
        per_vals = self.math.finalize(pdf) if m else torch.zeros((0, nq, 3), dtype=torch.float64, device=dev)
        return {
            "rows": (lo, hi), "pdf1d": pdf, "per_vals": per_vals,
            "exp_vals": small[:, :nq] / wsum[lo:hi, None], "best_vals": small[:, nq:2 * nq],
            "Pmax_indx": small[:, 2 * nq].to(torch.int64), "Pmax": top[lo:hi],
            "best_gridsub_tag": owner[lo:hi], "chi2min": chi2_all.min(0).values[lo:hi],
            "total_log_norm": merged_tln[lo:hi],
        }

    def _timed(self, fn):
        if not self.time_collectives:
            return fn()
        import torch

        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        self.collective_ms += e0.elapsed_time(e1)


def local_result_tensors(setup, res, device):
    """FitResult (host) -> the dict of device tensors ``merge_subgrids`` consumes; indices are
    mapped to rows of this subgrid as the reference stores them (fit.py:719-728)."""
    import torch

    g0 = setup.g0_indxs
    best_full = g0[np.maximum(res.Pmax_j, 0)]
    best_vals = np.stack([setup.qcols[k][best_full] for k in range(setup.n_q)], axis=1)
    host = {"total_log_norm": res.total_log_norm, "Pmax": res.Pmax, "Pmax_indx": best_full,
            "chi2min": res.chi2min, "best_vals": best_vals, "exp_vals": res.exp_vals,
            "pdf1d": res.pdf1d}
    return {k: torch.from_numpy(np.ascontiguousarray(v)).to(device) for k, v in host.items()}


class _BinLayout:
    """what CudaMergeMath needs of a FitSetup, rebuilt from a pdf1d file: the bin centres are
    the last row of every extension (fit.py:507-508)"""

    def __init__(self, bin_rows):
        self.n_q = len(bin_rows)
        self.bin_off = np.zeros(self.n_q + 1, dtype=np.int32)
        self.bin_off[1:] = np.cumsum([len(r) for r in bin_rows])
        self.total_bins = int(self.bin_off[-1])
        self.max_bins = int(max(len(r) for r in bin_rows))
        self.bin_val = np.concatenate(bin_rows).astype(np.float64)


def merge_pdf1d_stats(subgrid_pdf1d_fnames, subgrid_stats_fnames, re_run=False,
                      output_fname_base=None, partial=False, device=None, math=None):
    """Drop-in for ``beast.tools.subgridding_tools.merge_pdf1d_stats``
    (subgridding_tools.py:274-539): merge the 1D PDF and stats files of the subgrid fits of one
    catalog.  Same arguments, checks (``AssertionError`` on mismatching files), output files
    (``<base>_pdf1d.fits`` / ``<base>_stats.fits`` or ``combined_*``) and return value; the
    arithmetic runs on the GPU through ``merge_subgrids`` (``csrc/merge.cu``).

    ``device`` / ``math`` are additions: the CUDA device to use, or an object with the
    ``weights`` / ``scale_rows`` / ``finalize`` methods of ``CudaMergeMath``."""
    import os
    import re

    import torch

    from . import fitsio
    from .fit import save_pdf1d

    pdf1d_fname = (output_fname_base + "_pdf1d.fits") if output_fname_base is not None else "combined_pdf1d.fits"
    stats_fname = (output_fname_base + "_stats.fits") if output_fname_base is not None else "combined_stats.fits"
    if os.path.isfile(pdf1d_fname) and os.path.isfile(stats_fname) and re_run is False:
        print(str(len(subgrid_pdf1d_fnames)) + " files already merged, skipping")
        return pdf1d_fname, stats_fname
    nsub = len(subgrid_pdf1d_fnames)
    if not len(subgrid_stats_fnames) == nsub:
        raise AssertionError()

    def live_rows(ext_by_name):
        tot = np.sum(ext_by_name["M_ini"], axis=1)
        return len(np.where(tot > 0.0)[0]) - 1

    pdf_files = []
    for fname in subgrid_pdf1d_fnames:
        hdus = fitsio.read_fits(fname)[1:]
        pdf_files.append([(h["EXTNAME"], np.asarray(d)) for h, d in hdus])
    qnames = [q for q, _ in pdf_files[0]]
    first = dict(pdf_files[0])
    nobs = live_rows(first) if partial else first[qnames[0]].shape[0] - 1
    for other in pdf_files[1:]:
        cur = dict(other)
        for q in qnames:
            a0, a = first[q], cur[q]
            if not a0.shape[1] == a.shape[1]:
                raise AssertionError()
            if not partial:
                if not a0.shape[0] == a.shape[0]:
                    raise AssertionError()
            else:
                nobs = min(nobs, live_rows(cur))
            if not (np.isnan(a0[-1, 0]) and np.isnan(a[-1, 0]) or (a0[-1, :] == a[-1, :]).all()):
                raise AssertionError()

    stats, filters_tab = [], None
    for i, fname in enumerate(subgrid_stats_fnames):
        hdus = fitsio.read_fits(fname)
        stats.append(hdus[1][1])
        if i == 0 and len(hdus) > 2 and isinstance(hdus[2][1], dict):
            filters_tab = hdus[2][1]

    layout = _BinLayout([first[q][-1, :] for q in qnames])
    pcols = {}  # parameter -> percentiles present in the stats table, in column order
    for col in stats[0].keys():
        suffix = col.split("_")[-1]
        if re.compile(r"p\d{1,2}$").match(suffix):
            pcols.setdefault(col[: -len(suffix) - 1], []).append(int(suffix[1:]))
    pvals = sorted({p for v in pcols.values() for p in v}) or [16, 50, 84]
    trios = [(pvals[i:i + 3] + [50, 50])[:3] for i in range(0, len(pvals), 3)]
    if math is None:
        math = CudaMergeMath(layout, trios[0])
        dev = torch.device("cuda" if device is None else device)
    else:
        math.set_percentiles(trios[0])
        dev = torch.device("cpu" if device is None else device)

    def t(a, dtype=torch.float64):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev, dtype=dtype)

    parts = []
    for g in range(nsub):
        cur = dict(pdf_files[g])
        st = stats[g]
        parts.append({
            "total_log_norm": t(np.asarray(st["total_log_norm"])[:nobs]),
            "Pmax": t(np.asarray(st["Pmax"])[:nobs]),
            "Pmax_indx": t(np.asarray(st["Pmax_indx"])[:nobs], torch.int64),
            "chi2min": t(np.asarray(st["chi2min"])[:nobs]),
            "best_vals": t(np.stack([np.asarray(st[q + "_Best"])[:nobs] for q in qnames], axis=1)),
            "exp_vals": t(np.stack([np.asarray(st[q + "_Exp"])[:nobs] for q in qnames], axis=1)),
            "pdf1d": t(np.concatenate([cur[q][:nobs, :] for q in qnames], axis=1)),
        })
    merged = merge_subgrids(parts, math)
    host = {k: v.cpu().numpy() for k, v in merged.items()}
    per = {p: host["per_vals"][:, :, j] for j, p in enumerate(pvals[:3])}
    for i, trio in enumerate(trios[1:], start=1):  # the device routine takes three per pass
        math.set_percentiles(trio)
        block = math.finalize(merged["pdf1d"].clone()).cpu().numpy()
        for j, p in enumerate(pvals[3 * i:3 * i + 3]):
            per[p] = block[:, :, j]

    save_pdf1d_vals = []
    for k, q in enumerate(qnames):
        arr = np.zeros((nobs + 1, first[q].shape[1]))
        arr[:-1, :] = host["pdf1d"][:, layout.bin_off[k]:layout.bin_off[k + 1]]
        arr[-1, :] = first[q][-1, :]
        save_pdf1d_vals.append(arr)
    save_pdf1d(pdf1d_fname, save_pdf1d_vals, qnames)

    stats_dict = {}
    for col in stats[0].keys():
        suffix = col.split("_")[-1]
        if suffix == "Best" and col[:-5] in qnames:
            stats_dict[col] = host["best_vals"][:, qnames.index(col[:-5])]
        elif suffix == "Exp" and col[:-4] in qnames:
            stats_dict[col] = host["exp_vals"][:, qnames.index(col[:-4])]
        elif re.compile(r"p\d{1,2}$").match(suffix):
            stats_dict[col] = per[int(suffix[1:])][:, qnames.index(col[: -len(suffix) - 1])]
        elif col == "chi2min":
            stats_dict[col] = host["chi2min"]
        elif col == "Pmax":
            stats_dict[col] = host["Pmax"]
        elif col == "Pmax_indx":
            stats_dict[col] = host["Pmax_indx"]
        elif col == "total_log_norm":
            stats_dict[col] = host["total_log_norm"]
        elif not col == "chi2min_indx" and not col == "specgrid_indx":
            stats_dict[col] = np.asarray(stats[0][col])[:nobs]
    stats_dict["best_gridsub_tag"] = host["best_gridsub_tag"]
    fitsio.write_table_file(stats_fname, [stats_dict] + ([filters_tab] if filters_tab is not None else []))
    print("Saved combined 1dpdfs in " + pdf1d_fname)
    print("Saved combined stats in " + stats_fname)
    return pdf1d_fname, stats_fname


def merge_lnp(subgrid_lnp_fnames, re_run=False, output_fname_base=None, threshold=None):
    """Drop-in for ``beast.tools.subgridding_tools.merge_lnp`` (subgridding_tools.py:542-658):
    per star, concatenate the sparse (idx, lnp) of every subgrid file in the order given, tag
    each entry with the subgrid number taken from the ``gridsub<N>`` part of the file name,
    drop entries that are not within ``threshold`` of the star's merged maximum, and pad all
    stars to the same length with NaN (idx, subgrid) / -inf (lnp) (:629-656).  Like the
    reference's reader (tools/read_beast_data.py:47-57) the stars of one file are first padded
    to that file's longest list; with ``threshold=None`` those padding entries (lnp = -inf,
    idx = -1 here) stay in the output."""
    import os

    from .fit import read_lnp_groups, write_lnp_groups

    merged_lnp_fname = "combined_lnp.hd5" if output_fname_base is None else output_fname_base + "_lnp.hd5"
    if (os.path.isfile(merged_lnp_fname) or os.path.isfile(merged_lnp_fname + ".npz")) and re_run is False:
        print(str(len(subgrid_lnp_fnames)) + " files already merged, skipping")
        return merged_lnp_fname
    per_star, n_star = None, 0
    for fname in subgrid_lnp_fnames:
        subgrid_num = int([i for i in os.path.basename(fname).split("_") if "gridsub" in i][0][7:])
        groups = read_lnp_groups(fname)
        labels = [k for k in groups.keys() if "star" in k]
        n_star = len(labels)
        n_lnp = max(len(groups[k]["lnp"]) for k in labels)
        if per_star is None:
            per_star = [([], [], []) for _ in range(n_star)]
        for i in range(n_star):
            grp = groups["star_%d" % i]
            pad = n_lnp - len(grp["lnp"])
            per_star[i][0].append(np.concatenate([np.asarray(grp["idx"], dtype=float), np.full(pad, -1.0)]))
            per_star[i][1].append(np.concatenate([np.asarray(grp["lnp"], dtype=float), np.full(pad, -np.inf)]))
            per_star[i][2].append(np.full(n_lnp, float(subgrid_num)))
    merged = [merge_star_lnp(np.concatenate(s[0]), np.concatenate(s[1]), np.concatenate(s[2]), threshold)
              for s in per_star]
    longest = max(len(m[1]) for m in merged) if threshold is not None else 0
    groups = {}
    for i, (idx, lnp, sub) in enumerate(merged):
        pad = max(longest - len(lnp), 0)
        groups["star_%d" % i] = {
            "idx": np.concatenate([idx, np.full(pad, np.nan)]),
            "lnp": np.concatenate([lnp, np.full(pad, -np.inf)]),
            "subgrid": np.concatenate([sub, np.full(pad, np.nan)]),
        }
    write_lnp_groups(merged_lnp_fname, groups)
    return merged_lnp_fname


def merge_star_lnp(idx, lnp, sub, threshold):
    """one star of merge_lnp: keep the entries within ``threshold`` of the merged maximum
    (subgridding_tools.py:606-627); ``threshold=None`` keeps everything"""
    if threshold is not None and len(lnp):
        keep = np.where((lnp - lnp.max()) > threshold)[0]
        idx, lnp, sub = idx[keep], lnp[keep], sub[keep]
    return idx, lnp, sub


def merge_sparse_lnp(per_subgrid, threshold=-10.0):
    """merge_lnp (subgridding_tools.py:589-627): concatenate each star's (idx, lnp) over the
    subgrids, tag the subgrid, keep entries within ``threshold`` of the merged maximum.
    ``per_subgrid[g]`` = (sparse_off, idx, lnp) of subgrid g."""
    ns = len(per_subgrid[0][0]) - 1
    out = []
    for e in range(ns):
        idx = np.concatenate([p[1][p[0][e]:p[0][e + 1]] for p in per_subgrid])
        lnp = np.concatenate([p[2][p[0][e]:p[0][e + 1]] for p in per_subgrid])
        sub = np.concatenate([np.full(p[0][e + 1] - p[0][e], g) for g, p in enumerate(per_subgrid)])
        if threshold is not None and len(lnp):
            keep = np.where((lnp - lnp.max()) > threshold)[0]
            idx, lnp, sub = idx[keep], lnp[keep], sub[keep]
        out.append((idx, lnp, sub))
    return out
