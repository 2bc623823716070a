"""
Device-side driver of the fitting hot path: owns the resident grid (torch tensors as
plain device buffers) and runs the kernel pipeline of ``include/beast_b200.h`` on
batches of stars.

    dense FP32 sweep  ->  group selection  ->  exact FP64 max  ->  exact sparse pass
    (bf_dense_chi2)      (bf_select_*)        (bf_exact_max)      (bf_exact_accumulate)
                                                         [-> bf_pdf2d_accumulate, bf_sparse_gather]

This is the star loop of the reference (``beast/fitting/fit.py:637-762``) for a whole
batch at once.  PyTorch is used for allocation, streams and host<->device copies only.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import (GROUP, SELECT_SPLIT, STAR_BLOCK, AccumScratch, ExactGrid, ParamTables,
                    StarOutputs)


def _torch():
    import torch

    return torch


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class FitResult:
    """Per-star outputs of one ``FitEngine.fit`` call (host NumPy arrays, page-locked so the
    device->host copies run at full PCIe speed and asynchronously).

    Index arrays are in *filtered* model space (``j``); ``FitSetup.g0_indxs[j]`` maps to
    full-grid indices as the reference reports them (fit.py:719-728)."""

    _FIELDS = (("Pmax", "f8", ()), ("Pmax_j", "i8", ()), ("total_log_norm", "f8", ()),
               ("chi2min", "f8", ()), ("chi2min_j", "i8", ()), ("n_sparse", "i4", ()),
               ("n_groups_selected", "i8", ()))

    def __init__(self, ns, nq, total_bins, pair_len=0, sparse_cap=0):
        torch = _torch()
        self._t = {}
        self.ns = ns

        def alloc(name, shape, dtype):
            t = torch.empty(shape, dtype=dtype, pin_memory=True)
            self._t[name] = t
            setattr(self, name, t.numpy())

        for name, kind, _ in self._FIELDS:
            alloc(name, (ns,), {"f8": torch.float64, "i8": torch.int64, "i4": torch.int32}[kind])
        alloc("exp_vals", (ns, nq), torch.float64)
        alloc("per_vals", (ns, nq, 3), torch.float64)
        alloc("pdf1d", (ns, total_bins), torch.float64)
        self.pdf2d = None
        if pair_len:
            alloc("pdf2d", (ns, pair_len), torch.float64)
        self.sparse_off = None  # int64 [ns+1]
        self.sparse_j = self.sparse_lnp = self.sparse_chi2 = None
        self.sparse_cap = sparse_cap
        if sparse_cap:
            alloc("_sp_j", (sparse_cap,), torch.int64)
            alloc("_sp_lnp", (sparse_cap,), torch.float32)
            alloc("_sp_chi2", (sparse_cap,), torch.float32)


class FitEngine:
    """Grid resident in HBM + the batched fit pipeline.

    Parameters
    ----------
    setup : gridprep.FitSetup
    device : int or torch.device, default: current CUDA device
    gmin_budget_bytes : size cap of the (group x star) minimum map of one batch
    """

    def __init__(self, setup, device=None, gmin_budget_bytes=8 << 30, scratch_budget_bytes=12 << 30,
                 sparse_pass="split"):
        torch = _torch()
        self.lib = _cabi.load()
        if not torch.cuda.is_available():
            raise RuntimeError("beast_b200.FitEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("beast_b200.FitEngine: device must be a CUDA device")
        self.setup = setup
        sm, smem, khz = C.c_int(), C.c_int(), C.c_int()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _cabi.check(self.lib.bf_device_info(idx, C.byref(sm), C.byref(smem), C.byref(khz)), "bf_device_info")
        self.sm_count, self.clock_khz = sm.value, khz.value

        def up(a):
            return torch.from_numpy(np.ascontiguousarray(a)).to(self.device)

        s = setup
        with torch.cuda.device(self.device):
            self.d_tiles = up(s.tiles)
            self.d_tiles_coarse = up(s.tiles_coarse)
            self.d_msb, self.d_ivar = up(s.msb), up(s.ivar)
            self.d_lnq, self.d_lnw = up(s.lnq), up(s.lnw)
            self.d_cov_off = up(s.cov_off) if s.cov_off is not None else None
            split_only = sparse_pass == "split" and s.split_ok and not s.full_cov and s.nb <= 8
            # (the per-model tables are only read by the kernels that treat all parameters alike)
            self.d_bin_tab, self.d_q_tab = (None, None) if split_only else (up(s.bin_tab), up(s.q_tab))
            self.d_bin_off, self.d_bin_val = up(s.bin_off), up(s.bin_val)
            self.d_q_lin = up(s.q_lin)
            self.d_q_const, self.d_bin_q = up(s.q_const), up(s.bin_q)
            # sparse_pass: "split" = group-ordered weights + class-split histogram kernel (needs
            # the class tables; falls back when the grid has more than 32 parameters per class),
            # "legacy" = star-ordered weights + full-width / windowed histogram kernels,
            # "both" = all tables resident, the kernels chosen per call (tests)
            if sparse_pass not in ("split", "legacy", "both"):
                raise ValueError("sparse_pass must be 'split', 'legacy' or 'both'")
            # (the group-ordered weights kernel: diagonal noise model, up to 8 bands)
            self.split = sparse_pass != "legacy" and s.split_ok and not s.full_cov and s.nb <= 8
            self.sparse_pass = sparse_pass
            if self.split:
                self.d_fast_k, self.d_slow_k = up(s.fast_k), up(s.slow_k)
                self.d_fast_bin, self.d_fast_q = up(s.fast_bin), up(s.fast_q)
                self.d_slow_bin, self.d_slow_q = up(s.slow_bin), up(s.slow_q)
                self.d_slow_run = up(s.slow_run)
            self.d_inv_scale = up((1.0 / s.band_scale).astype(np.float64))
            if s.pair_bin is not None:
                self.d_pair_bin, self.d_pair_off = up(s.pair_bin), up(s.pair_off)
            else:
                self.d_pair_bin = self.d_pair_off = None
        self.grid = ExactGrid(s.nb, s.n_models, s.ld_model, self.d_msb.data_ptr(),
                              self.d_ivar.data_ptr(), self.d_lnq.data_ptr(), self.d_lnw.data_ptr(),
                              self.d_cov_off.data_ptr() if self.d_cov_off is not None else None)
        dp = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
        self.params = ParamTables(s.n_q, s.total_bins, s.max_bins, dp(self.d_bin_tab),
                                  dp(self.d_q_tab), self.d_q_lin.data_ptr(),
                                  self.d_bin_off.data_ptr(), self.d_bin_val.data_ptr(),
                                  self.d_q_const.data_ptr(), self.d_bin_q.data_ptr())
        if self.split:
            p = self.params
            p.n_fast, p.n_slow = len(s.fast_k), len(s.slow_k)
            p.fast_pad, p.slow_pad = s.fast_pad, s.slow_pad
            p.fast_k, p.slow_k = self.d_fast_k.data_ptr(), self.d_slow_k.data_ptr()
            p.fast_bin, p.fast_q = self.d_fast_bin.data_ptr(), self.d_fast_q.data_ptr()
            p.slow_bin, p.slow_q = self.d_slow_bin.data_ptr(), self.d_slow_q.data_ptr()
            p.slow_run = self.d_slow_run.data_ptr()
        per_star = s.n_groups * 4
        mb = max(STAR_BLOCK, (gmin_budget_bytes // per_star) // STAR_BLOCK * STAR_BLOCK)
        mb = min(mb, ((2**31 - 1) // max(s.n_groups, 1)) // STAR_BLOCK * STAR_BLOCK, 32768 * STAR_BLOCK)
        self.max_batch = int(max(mb, STAR_BLOCK))
        # per listed group: 64 weights + partials (+ the group order and the chunk lists)
        per_entry = GROUP * 8 + 24 + (4 + 128 + 64 + 8 * s.fast_pad if self.split else 0)
        self.max_entries = int(max(1 << 16, scratch_budget_bytes // per_entry))
        self.launches = 0  # kernels launched by this engine (bench.py reports it)
        self._result_cache = {}
        self._copy_stream = torch.cuda.Stream(device=self.device)
        self._aux_stream = torch.cuda.Stream(device=self.device)
        self.profile = False  # when True, CUDA events bracket every stage (see stage_ms())
        self._events = []

    # ------------------------------------------------------------------------------
    class _Timed:
        def __init__(self, eng, name):
            self.eng, self.name = eng, name

        def __enter__(self):
            if self.eng.profile:
                torch = _torch()
                self.e0 = torch.cuda.Event(enable_timing=True)
                self.e1 = torch.cuda.Event(enable_timing=True)
                self.e0.record(torch.cuda.current_stream(self.eng.device))

        def __exit__(self, *exc):
            if self.eng.profile:
                self.e1.record(_torch().cuda.current_stream(self.eng.device))
                self.eng._events.append((self.name, self.e0, self.e1))

    def _timed(self, name):
        return FitEngine._Timed(self, name)

    def stage_ms(self, reset=True):
        """per-stage device time (ms, summed over the launches since the last reset) and launch
        counts, from CUDA events on the launching stream; call after a synchronize."""
        tot, cnt = {}, {}
        for name, e0, e1 in self._events:
            tot[name] = tot.get(name, 0.0) + e0.elapsed_time(e1)
            cnt[name] = cnt.get(name, 0) + 1
        if reset:
            self._events = []
        return tot, cnt

    # ------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(_torch().cuda.current_stream(self.device).cuda_stream)

    def _flux32(self, d_flux64, n):
        torch = _torch()
        ld = (n + STAR_BLOCK - 1) // STAR_BLOCK * STAR_BLOCK
        f32 = torch.zeros((self.setup.nb, ld), dtype=torch.float32, device=self.device)
        f32[:, :n] = (d_flux64 * self.d_inv_scale[None, :]).t().to(torch.float32)
        return f32, ld

    def dense(self, d_flux64, n, coarse=False):
        """dense sweep of one batch over the full grid (or its coarse subsample);
        returns (gmin, starmin, ld_star)"""
        torch = _torch()
        s = self.setup
        f32, ld = self._flux32(d_flux64, n)
        tiles = self.d_tiles_coarse if coarse else self.d_tiles
        n_tiles = int(tiles.shape[0])
        gmin = torch.empty((n_tiles * (_cabi.TILE_MODELS // GROUP), ld), dtype=torch.float32,
                           device=self.device)
        # packed (FP32 bits << 32 | arg-min group), start at (+inf, no group)
        starmin = torch.full((ld,), 0x7F800000FFFFFFFF, dtype=torch.int64, device=self.device)
        tpi = self.lib.bf_dense_tiles_per_item(n_tiles, n, self.sm_count)
        with self._timed("presort" if coarse else "dense"):
            _cabi.call("bf_dense_chi2_cov" if s.full_cov else "bf_dense_chi2", s.nb, n_tiles,
                       _ptr(tiles), n, ld, _ptr(f32),
                       _ptr(gmin), _ptr(starmin), tpi, self._stream())
        self.launches += 1
        return gmin, starmin, ld

    def select(self, gmin, starmin, ld, n, threshold):
        """group lists of one batch; returns (offsets[n*SPLIT+1], list)"""
        torch = _torch()
        s = self.setup
        tau = float(-2.0 * threshold)
        nc = n * SELECT_SPLIT
        counts = torch.empty(nc, dtype=torch.int32, device=self.device)
        offsets = torch.empty(nc + 1, dtype=torch.int32, device=self.device)
        scratch = torch.empty(int(self.lib.bf_scan_scratch_elems(nc)), dtype=torch.int32, device=self.device)
        st = self._stream()
        with self._timed("select_count"):
            _cabi.call("bf_select_count", s.n_groups, n, ld, _ptr(gmin), _ptr(starmin), tau,
                       s.err_a, s.err_b, _ptr(counts), st)
        _cabi.call("bf_exclusive_scan_u32", nc, _ptr(counts), _ptr(offsets), _ptr(scratch), st)
        total = int(offsets[-1].item()) & 0xFFFFFFFF  # the one host sync of a batch
        lst = torch.empty(max(total, 1), dtype=torch.int32, device=self.device)
        entry_star = torch.empty(max(total, 1), dtype=torch.int32, device=self.device)
        with self._timed("select_fill"):
            _cabi.call("bf_select_fill", s.n_groups, n, ld, _ptr(gmin), _ptr(starmin), tau,
                       s.err_a, s.err_b, _ptr(offsets), _ptr(lst), _ptr(entry_star), st)
        self.launches += 5
        return offsets, lst, entry_star, total

    def _sparse_records(self, o, sparse, lnp_rng, rng, seed, star_base, d_flux, offsets, lst, mask,
                        prefix, n):
        """sparse lnP records (fit.py:693-709).  ``sparse`` = "all" or lnp_npts.

        lnp_rng="numpy": ranks drawn on the host with np.random.choice(n, size, replace=False),
        the call the reference makes (only the set size feeds its random stream), so a seeded
        run reproduces the reference's selection.  lnp_rng="device": ranks from
        bf_sparse_sample_ranks (same distribution, O(lnp_npts) per star, no host round trip
        of the sets)."""
        torch = _torch()
        dev, st = self.device, self._stream()
        n_sp = o["n_sparse"].to(torch.int64) & 0xFFFFFFFF
        ranks = None
        if sparse == "all" or lnp_rng == "device":
            cnt = n_sp if sparse == "all" else torch.clamp(n_sp, max=int(sparse))
            d_rec_off = torch.zeros(n + 1, dtype=torch.int64, device=dev)
            d_rec_off[1:] = torch.cumsum(cnt, 0)
            nrec = int(d_rec_off[-1].item())
            if sparse != "all" and nrec:
                ranks = torch.empty(nrec, dtype=torch.int32, device=dev)
                _cabi.call("bf_sparse_sample_ranks", n, _ptr(o["n_sparse"]), _ptr(d_rec_off),
                           int(seed) & 0xFFFFFFFFFFFFFFFF, int(star_base), _ptr(ranks), st)
                self.launches += 1
        else:
            ns_host = n_sp.cpu().numpy()
            npts = int(sparse)
            choose = rng.choice if rng is not None else np.random.choice
            parts = [np.asarray(choose(int(m), size=npts, replace=False)) if npts < m else np.arange(m)
                     for m in ns_host]
            rec_off = np.zeros(n + 1, dtype=np.int64)
            rec_off[1:] = np.cumsum(np.minimum(ns_host, npts))
            nrec = int(rec_off[-1])
            d_rec_off = torch.from_numpy(rec_off).to(dev)
            if nrec:
                ranks = torch.from_numpy(np.concatenate(parts).astype(np.int32)).to(dev)
        out_j = torch.empty(max(nrec, 1), dtype=torch.int64, device=dev)
        out_lnp = torch.empty(max(nrec, 1), dtype=torch.float32, device=dev)
        out_chi2 = torch.empty(max(nrec, 1), dtype=torch.float32, device=dev)
        if nrec:
            with self._timed("sparse_gather"):
                _cabi.call("bf_sparse_gather", C.byref(self.grid), n, _ptr(d_flux), _ptr(offsets),
                           _ptr(lst), _ptr(mask), _ptr(prefix), _ptr(d_rec_off), _ptr(ranks),
                           _ptr(out_j), _ptr(out_lnp), _ptr(out_chi2), st)
            self.launches += 1
        o["sparse_off"] = d_rec_off
        o["sparse_j"], o["sparse_lnp"], o["sparse_chi2"] = out_j[:nrec], out_lnp[:nrec], out_chi2[:nrec]

    def fit_batch(self, flux, threshold=-10.0, percentiles=(16.0, 50.0, 84.0), want_pdf2d=False,
                  sparse=None, rng=None, lnp_rng="numpy", seed=0, star_base=0, presort=None):
        """Fit one batch (<= max_batch stars).  ``flux``: (n, nb) float64 NumPy array or
        torch tensor (host or device).  Returns a dict of device tensors in input order.

        presort (default: on, except with lnp_rng="numpy" subsampling, whose random stream
        follows catalog order): the batch is first swept against every 32nd model, stars are
        ordered by the arg-min of that coarse sweep, and the whole pipeline runs in that order
        -- neighbouring stars then select the same groups, so the exact pass finds its grid
        and table lines in L2.  Outputs are put back in input order at the end."""
        torch = _torch()
        s = self.setup
        with torch.cuda.device(self.device):
            if isinstance(flux, np.ndarray):
                flux = torch.from_numpy(np.ascontiguousarray(flux, dtype=np.float64))
            d_flux = flux.to(self.device, dtype=torch.float64, non_blocking=True).contiguous()
            n = d_flux.shape[0]
            if n > self.max_batch:
                raise ValueError("batch of %d stars exceeds max_batch=%d" % (n, self.max_batch))
            if presort is None:
                presort = not (sparse not in (None, "all") and lnp_rng == "numpy")
            perm = None
            if presort and n > 1:
                _, cmin, _ = self.dense(d_flux, n, coarse=True)
                perm = torch.argsort(cmin[:n] & 0xFFFFFFFF)
                d_flux = d_flux[perm].contiguous()
            gmin, starmin, ld = self.dense(d_flux, n)
            offsets, lst, entry_star, total = self.select(gmin, starmin, ld, n, threshold)
            st = self._stream()
            dev = self.device
            pmax = torch.empty(n, dtype=torch.float64, device=dev)
            pmax_j = torch.empty(n, dtype=torch.int64, device=dev)
            with self._timed("exact_max"):
                _cabi.call("bf_exact_max", C.byref(self.grid), n, ld, _ptr(d_flux), _ptr(gmin),
                           _ptr(starmin), s.err_a, s.err_b, _ptr(offsets), _ptr(lst), _ptr(pmax),
                           _ptr(pmax_j), st)
            del gmin
            o = {
                "Pmax": pmax, "Pmax_j": pmax_j,
                "total_log_norm": torch.empty(n, dtype=torch.float64, device=dev),
                "chi2min": torch.empty(n, dtype=torch.float64, device=dev),
                "chi2min_j": torch.empty(n, dtype=torch.int64, device=dev),
                "n_sparse": torch.empty(n, dtype=torch.int32, device=dev),
                "exp_vals": torch.empty((n, s.n_q), dtype=torch.float64, device=dev),
                "pdf1d": torch.empty((n, s.total_bins), dtype=torch.float64, device=dev),
                "per_vals": torch.empty((n, s.n_q, 3), dtype=torch.float64, device=dev),
            }
            mask = torch.empty(max(total, 1) * (GROUP // 32), dtype=torch.int32, device=dev)
            prefix = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
            pct = (C.c_double * 3)(*[float(x) for x in percentiles])
            off_star = (offsets.view(-1)[:: SELECT_SPLIT].to(torch.int64) & 0xFFFFFFFF)
            off_host = off_star.cpu().numpy()
            # sub-ranges of stars whose lists fit the scratch budget
            cuts, a0 = [], 0
            while a0 < n:
                b0 = int(np.searchsorted(off_host, off_host[a0] + self.max_entries, side="right")) - 1
                b0 = min(n, max(b0, a0 + 1))
                cuts.append((a0, b0))
                a0 = b0
            nmax = max(int(off_host[b0] - off_host[a0]) for a0, b0 in cuts)
            w_scr = torch.empty(max(nmax, 1) * GROUP, dtype=torch.float64, device=dev)
            p_w = torch.empty(max(nmax, 1), dtype=torch.float64, device=dev)
            p_c = torch.empty(max(nmax, 1), dtype=torch.float64, device=dev)
            p_j = torch.empty(max(nmax, 1), dtype=torch.int64, device=dev)
            flags = torch.empty(max(b - a for a, b in cuts), dtype=torch.int32, device=dev)
            scr = AccumScratch(w_scr.data_ptr(), p_w.data_ptr(), p_c.data_ptr(), p_j.data_ptr(),
                               flags.data_ptr())
            use_split = self.split and not getattr(self, "force_legacy", False)
            if use_split:
                g_order = torch.empty(max(nmax, 1), dtype=torch.int32, device=dev)
                c_refs = torch.empty(max(nmax, 1) * 32, dtype=torch.int32, device=dev)
                g_work = torch.empty(int(self.lib.bf_group_order_work_elems(s.n_groups)),
                                     dtype=torch.int32, device=dev)
                c_wsum = torch.empty(max(nmax, 1) * 8, dtype=torch.float64, device=dev)
                e_part = torch.empty(max(nmax, 1) * s.fast_pad, dtype=torch.float64, device=dev)
                scr.group_order, scr.chunk_refs = g_order.data_ptr(), c_refs.data_ptr()
                scr.chunk_wsum, scr.exp_part = c_wsum.data_ptr(), e_part.data_ptr()
            outs = StarOutputs(o["total_log_norm"].data_ptr(), o["chi2min"].data_ptr(),
                               o["chi2min_j"].data_ptr(), o["n_sparse"].data_ptr(),
                               o["exp_vals"].data_ptr(), o["pdf1d"].data_ptr(),
                               o["per_vals"].data_ptr(), mask.data_ptr(), prefix.data_ptr())
            # (running the weights kernel of the next range beside the histogram kernel of this one
            # was measured and is slower: the latency-bound histogram kernel stalls once the
            # memory-bound kernels saturate L2 -- see DESIGN.md)
            n_listed = off_star[1:] - off_star[:-1]
            pdf2d = None
            if want_pdf2d and self.d_pair_bin is not None:
                pdf2d = torch.zeros((n, int(s.pair_off[-1])), dtype=torch.float64, device=dev)
            reeval = pdf2d is not None and getattr(self, "pdf2d_reevaluate", False)
            for a0, b0 in cuts:
                base, cnt = int(off_host[a0]), int(off_host[b0] - off_host[a0])
                # one block per star in the histogram kernel: longest lists first, so that the
                # last blocks of the launch are short ones
                order = torch.argsort(n_listed[a0:b0], descending=True, stable=True).to(torch.int32)
                if use_split and cnt:
                    with self._timed("group_order"):
                        _cabi.call("bf_group_order", s.n_groups, _ptr(lst), base, cnt, _ptr(g_work),
                                   _ptr(g_order), st)
                    self.launches += 5
                with self._timed("accumulate"):
                    _cabi.call("bf_exact_accumulate", C.byref(self.grid), C.byref(self.params),
                               a0, b0 - a0, _ptr(d_flux), _ptr(pmax), float(threshold),
                               _ptr(offsets), _ptr(lst), _ptr(entry_star), _ptr(order), base, cnt,
                               pct, C.byref(scr), C.byref(outs), st,
                               C.c_void_p(self._aux_stream.cuda_stream))
                self.launches += 2 if use_split else 3
                if pdf2d is not None and cnt and not reeval:
                    # 2D PDFs of this range from the weights still in the scratch (fit.py:758-762)
                    with self._timed("pdf2d"):
                        _cabi.call("bf_pdf2d_from_weights", s.ld_model, len(s.pairs),
                                   _ptr(self.d_pair_bin), _ptr(self.d_pair_off), _ptr(pmax),
                                   _ptr(o["total_log_norm"]), _ptr(lst), _ptr(entry_star), _ptr(mask),
                                   _ptr(w_scr), base, cnt, _ptr(pdf2d), st)
                    self.launches += 1
            o["n_groups_selected"] = off_star[1:] - off_star[:-1]
            if reeval:  # the stand-alone entry point: evaluates the listed models again (tests)
                _cabi.call("bf_pdf2d_accumulate", C.byref(self.grid), len(s.pairs),
                           _ptr(self.d_pair_bin), _ptr(self.d_pair_off), n, _ptr(d_flux),
                           _ptr(pmax), _ptr(o["total_log_norm"]), float(threshold),
                           _ptr(offsets), _ptr(lst), _ptr(pdf2d), st)
                self.launches += 1
            if pdf2d is not None:
                o["pdf2d"] = pdf2d

            if sparse is not None:
                self._sparse_records(o, sparse, lnp_rng, rng, seed, star_base, d_flux, offsets, lst,
                                     mask, prefix, n)
            if perm is not None:
                self._unpermute(o, perm, n)
            return o

    def _unpermute(self, o, perm, n):
        """put the outputs of a presorted batch back in input order (row i of the sorted run
        belongs to input star perm[i])"""
        torch = _torch()
        for name in ("Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse",
                     "exp_vals", "pdf1d", "per_vals", "n_groups_selected", "pdf2d"):
            if name in o:
                dst = torch.empty_like(o[name])
                dst[perm] = o[name]
                o[name] = dst
        if "sparse_off" in o:
            off_s = o["sparse_off"]
            cnt_s = off_s[1:] - off_s[:-1]
            inv = torch.empty_like(perm)
            inv[perm] = torch.arange(n, device=perm.device)
            cnt = cnt_s[inv]
            off = torch.zeros(n + 1, dtype=torch.int64, device=perm.device)
            off[1:] = torch.cumsum(cnt, 0)
            nrec = int(o["sparse_j"].shape[0])
            if nrec:
                src = torch.repeat_interleave(off_s[:-1][inv] - off[:-1], cnt) + torch.arange(
                    nrec, device=perm.device)
                for name in ("sparse_j", "sparse_lnp", "sparse_chi2"):
                    o[name] = o[name][src]
            o["sparse_off"] = off

    # ------------------------------------------------------------------------------
    def fit(self, fluxes, threshold=-10.0, percentiles=(16.0, 50.0, 84.0), want_pdf2d=False,
            sparse=None, batch=None, rng=None, lnp_rng="numpy", seed=0, reuse_buffers=False):
        """Fit all stars of ``fluxes`` (ns, nb) in batches; returns a host ``FitResult``.

        Results are copied device->host asynchronously into page-locked arrays while the next
        batch computes.  With ``reuse_buffers=True`` the engine keeps those arrays and hands
        the same ones back on the next call of the same shape (they are overwritten then)."""
        torch = _torch()
        s = self.setup
        fluxes = np.ascontiguousarray(fluxes, dtype=np.float64)
        ns = fluxes.shape[0]
        if fluxes.ndim != 2 or fluxes.shape[1] != s.nb:
            raise ValueError("fluxes must have shape (n_stars, %d)" % s.nb)
        bs = min(batch or self.max_batch, self.max_batch)
        pair_len = int(s.pair_off[-1]) if (want_pdf2d and s.pair_off is not None) else 0
        if pair_len:  # 2D PDFs are dense per star: keep one batch of them within ~16 GB of HBM
            bs = int(max(1, min(bs, (16 << 30) // (pair_len * 8))))
        cap = ns * int(sparse) if (sparse is not None and sparse != "all") else 0
        key = (ns, pair_len, cap)
        res = self._result_cache.get(key) if reuse_buffers else None
        if res is None:
            res = FitResult(ns, s.n_q, s.total_bins, pair_len, cap)
            if reuse_buffers:
                self._result_cache = {key: res}
        t = res._t
        sp_parts, sp_cnt, keep = [], [], []
        sp_pos = 0
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream(self.device)
            copy_stream = self._copy_stream
            for a in range(0, ns, bs):
                b = min(ns, a + bs)
                o = self.fit_batch(fluxes[a:b], threshold, percentiles, want_pdf2d, sparse, rng,
                                   lnp_rng, seed, a)
                done = torch.cuda.Event()
                done.record(main)
                copy_stream.wait_event(done)
                with torch.cuda.stream(copy_stream):
                    names = ["Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse",
                             "n_groups_selected", "exp_vals", "per_vals", "pdf1d"]
                    if pair_len:
                        names.append("pdf2d")
                    for name in names:
                        o[name].record_stream(copy_stream)
                        t[name][a:b].copy_(o[name], non_blocking=True)
                    if sparse is not None:
                        off = o["sparse_off"]
                        nrec = int(o["sparse_j"].shape[0])
                        sp_cnt.append(off)
                        if cap:
                            for name, src in (("_sp_j", "sparse_j"), ("_sp_lnp", "sparse_lnp"),
                                              ("_sp_chi2", "sparse_chi2")):
                                o[src].record_stream(copy_stream)
                                t[name][sp_pos:sp_pos + nrec].copy_(o[src], non_blocking=True)
                            sp_pos += nrec
                        else:
                            sp_parts.append([o["sparse_j"].cpu(), o["sparse_lnp"].cpu(),
                                             o["sparse_chi2"].cpu()])
                keep.append(o)
            copy_stream.synchronize()
            main.synchronize()
        if sparse is not None:
            res.sparse_off = np.zeros(ns + 1, dtype=np.int64)
            res.sparse_off[1:] = np.cumsum(np.concatenate([np.diff(x.cpu().numpy()) for x in sp_cnt]))
            if cap:
                n = int(res.sparse_off[-1])
                res.sparse_j, res.sparse_lnp, res.sparse_chi2 = res._sp_j[:n], res._sp_lnp[:n], res._sp_chi2[:n]
            else:
                res.sparse_j = np.concatenate([p[0].numpy() for p in sp_parts])
                res.sparse_lnp = np.concatenate([p[1].numpy() for p in sp_parts])
                res.sparse_chi2 = np.concatenate([p[2].numpy() for p in sp_parts])
        return res
