"""
Drop-in for ``beast.fitting.trim_grid.trim_models`` (reference ``beast/fitting/trim_grid.py:10-147``):
drop the models that can never match the observed flux range, on the GPU.

``trim_indices`` returns the reference's ``indxs`` (ascending full-grid indices of the kept
models); ``trim_models`` keeps the reference's signature, builds the trimmed grid / noise model
(with the ``fullgrid_idx`` column, :125) and writes them with the reference's own classes when
they are importable (PyTables / h5py / astropy are not on every box); it also returns the trimmed
objects so that the fit can follow without a file round trip.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _cabi

COMPLETE_BIT = 1 << 31


def _data_range(obsdata, inFlux=True):
    """faintest / brightest observed flux per band (trim_grid.py:54-72)"""
    nb = len(obsdata.filters)
    lo, hi = np.zeros(nb), np.zeros(nb)
    for k, f in enumerate(obsdata.filters):
        col = np.asarray(obsdata.data[obsdata.filter_aliases[f]])
        vals = col * obsdata.vega_flux[k] if inFlux else 10 ** (-0.4 * col) * obsdata.vega_flux[k]
        lo[k], hi[k] = np.amin(vals), np.amax(vals)
    return lo, hi


def resolve_required(hist, nb):
    """The reference applies the band tests one after another and skips a band whose test would
    leave nothing (trim_grid.py:98-111).  ``hist[p | complete << nb]`` = number of models with
    pass pattern p.  Returns (required bit mask, n complete models, n kept)."""
    hist = np.asarray(hist, dtype=np.int64)
    pats = np.arange(len(hist), dtype=np.int64)
    complete = (pats >> nb) & 1
    n_ast = int(hist[complete == 1].sum())
    if n_ast <= 0:
        raise ValueError("no models are brighter than the minimum ASTs run")
    alive = complete == 1
    required = 0
    for k in range(nb):
        trial = alive & (((pats >> k) & 1) == 1)
        if hist[trial].sum() > 0:
            alive = trial
            required |= 1 << k
    n_keep = int(hist[alive].sum())
    if n_keep == 0:
        raise ValueError("no models that are within the data range")
    return required | COMPLETE_BIT, n_ast, n_keep


def trim_indices(seds, noisemodel, min_data, max_data, sigma_fac=3.0, device=None):
    """Indices of the models ``trim_models`` keeps (the reference's ``indxs``), ascending."""
    import torch

    lib = _cabi.load()
    if not torch.cuda.is_available():
        raise RuntimeError("beast_b200.trim needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    seds = np.ascontiguousarray(seds, dtype=np.float64)
    nm, nb = seds.shape
    if nb > 15:
        raise NotImplementedError("trim kernels support at most 15 bands")

    def up(a):
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)

    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        d_seds, d_bias = up(seds), up(noisemodel["bias"])
        d_unc, d_compl = up(noisemodel["error"]), up(noisemodel["completeness"])
        d_lo, d_hi = up(min_data), up(max_data)
        pattern = torch.empty(nm, dtype=torch.int32, device=dev)
        hist = torch.zeros(1 << (nb + 1), dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        _cabi.call("bf_trim_flags", nb, nm, p(d_seds), p(d_bias), p(d_unc), p(d_compl), float(sigma_fac),
                   p(d_lo), p(d_hi), p(pattern), p(hist), st)
        required, n_ast, n_keep = resolve_required(hist.cpu().numpy(), nb)
        keep = torch.empty(nm, dtype=torch.int32, device=dev)
        offsets = torch.empty(nm + 1, dtype=torch.int32, device=dev)
        scratch = torch.empty(int(lib.bf_scan_scratch_elems(nm)), dtype=torch.int32, device=dev)
        _cabi.call("bf_trim_select", nm, p(pattern), C.c_uint32(required & 0xFFFFFFFF), p(keep), st)
        _cabi.call("bf_exclusive_scan_u32", nm, p(keep), p(offsets), p(scratch), st)
        indices = torch.empty(n_keep, dtype=torch.int64, device=dev)
        _cabi.call("bf_trim_compact", nm, p(keep), p(offsets), p(indices), st)
        out = indices.cpu().numpy()
    return out, n_ast


def trim_models(sedgrid, sedgrid_noisemodel, obsdata, sed_outname=None, noisemodel_outname=None,
                sigma_fac=3.0, n_detected=4, inFlux=True, trunchen=False, device=None):
    """Same arguments as the reference (``n_detected`` is accepted and, as there, unused).

    Returns ``(trimmed_grid_columns, trimmed_seds, trimmed_noisemodel, indxs)``; files are
    written only when the reference's writers are importable."""
    min_data, max_data = _data_range(obsdata, inFlux)
    seds = sedgrid.seds.read() if hasattr(sedgrid.seds, "read") else np.asarray(sedgrid.seds)
    print("number of original models = ", len(seds))
    indxs, n_ast = trim_indices(seds, sedgrid_noisemodel, min_data, max_data, sigma_fac, device)
    print("number of ast trimmed models = ", n_ast)
    print("number of trimmed models = ", len(indxs))
    table = sedgrid.grid if hasattr(sedgrid, "grid") else sedgrid
    cols = {key: np.asarray(table[key])[indxs] for key in list(table.keys())}
    cols["fullgrid_idx"] = indxs.astype(int)
    names = ["bias", "error", "completeness"] + (["q_norm", "icov_diag", "icov_offdiag"] if trunchen else [])
    noise = {k: np.asarray(sedgrid_noisemodel[k])[indxs] for k in names}
    tseds = seds[indxs]
    if sed_outname is not None or noisemodel_outname is not None:
        try:
            import tables
            from astropy.table import Table
            from beast.physicsmodel.grid import SEDGrid
        except Exception as exc:
            raise RuntimeError("writing trimmed grid files needs the reference's SEDGrid, astropy and "
                               "PyTables, which are not importable here; call without output names "
                               "to get the trimmed arrays in memory") from exc
        g = SEDGrid(sedgrid.lamb, seds=tseds, grid=Table(cols), backend="memory")
        g.header["filters"] = " ".join(obsdata.filters)
        g.write(sed_outname)
        with tables.open_file(noisemodel_outname, "w") as outfile:
            for k in names:
                outfile.create_array(outfile.root, k, noise[k])
    return cols, tseds, noise, indxs
