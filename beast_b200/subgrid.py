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
        self.pct = (C.c_double * 3)(*[float(x) for x in percentiles])
        self._tables = None

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
