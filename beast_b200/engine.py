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

    def __init__(self, ns, nq, total_bins, pair_len=0, sparse_cap=0, extra_rows=0, extra_rows_2d=0):
        torch = _torch()
        self._t = {}
        self.ns = ns
        self.sparse_filled = 0

        def alloc(name, shape, dtype):
            t = torch.empty(shape, dtype=dtype, pin_memory=True)
            self._t[name] = t
            setattr(self, name, t.numpy())

        for name, kind, _ in self._FIELDS:
            alloc(name, (ns,), {"f8": torch.float64, "i8": torch.int64, "i4": torch.int32}[kind])
        alloc("exp_vals", (ns, nq), torch.float64)
        alloc("per_vals", (ns, nq, 3), torch.float64)
        # (extra rows: room for the bin-centre rows the writers append, fit.py:507-508, 569-573)
        alloc("pdf1d", (ns + extra_rows, total_bins), torch.float64)
        self.pdf2d = None
        if pair_len:
            alloc("pdf2d", (ns + extra_rows_2d, pair_len), torch.float64)
        self.sparse_off = None  # int64 [ns+1]
        self.sparse_j = self.sparse_lnp = self.sparse_chi2 = None
        self.sparse_cap = sparse_cap
        if sparse_cap:
            alloc("_sp_j", (sparse_cap,), torch.int64)
            alloc("_sp_lnp", (sparse_cap,), torch.float32)
            alloc("_sp_chi2", (sparse_cap,), torch.float32)


def bind_host_to_gpu(device_index):
    """Pin the calling process to the CPU cores next to GPU ``device_index`` (NVML's ideal CPU
    affinity), so that the page-locked result arrays allocated afterwards sit in that socket's
    memory and the device-to-host copies do not cross the socket link.  One process per GPU;
    call before the first allocation.  Returns the cores chosen, or None when NVML or the
    affinity calls are not available (nothing is changed then)."""
    import os

    try:
        import pynvml

        torch = _torch()
        pr = torch.cuda.get_device_properties(device_index)
        pynvml.nvmlInit()
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode() if bytes is not str else bus)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {i * 64 + b for i, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


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
            self.d_boxes = up(s.boxes) if s.boxes is not None else None
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
                self.d_fast_rng = up(s.fast_rng.view(np.int16))
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
            p.fast_rng = self.d_fast_rng.data_ptr()
        per_star = s.n_groups * 4
        mb = max(STAR_BLOCK, (gmin_budget_bytes // per_star) // STAR_BLOCK * STAR_BLOCK)
        mb = min(mb, ((2**31 - 1) // max(s.n_groups, 1)) // STAR_BLOCK * STAR_BLOCK, 32768 * STAR_BLOCK)
        self.max_batch = int(max(mb, STAR_BLOCK))
        # per listed group: 64 weights + partials (+ the group order and the chunk lists)
        self.scratch_budget_bytes = int(scratch_budget_bytes)
        self.presort = True          # coarse sweep + star ordering inside bf_fit_batch
        self.prune = True            # the dense sweep skips groups no star of a warp can select
        self.sort_band = setup.sort_band if not setup.full_cov else -1   # -1: by coarse arg-min group
        self.force_legacy = False    # True: never the class-split kernels (tests, measurements)
        self.launches = 0            # kernels launched by this engine (bench.py reports it)
        self._result_cache = {}
        self._workspace = None
        self._copy_stream = torch.cuda.Stream(device=self.device)
        self._aux_stream = torch.cuda.Stream(device=self.device)
        self.profile = False         # True: bf_fit_batch times its stages (synchronising)
        self.reuse_buffers = False   # True: Q_all_memory hands out this engine's page-locked arrays
        self._stage_tot, self._stage_cnt = {}, {}
        self.last_state = None

    # ------------------------------------------------------------------------------
    def _plan(self, want_pdf2d):
        s = self.setup
        p = _cabi.FitPlan()
        p.nb, p.full_cov, p.n_groups = s.nb, int(s.full_cov), s.n_groups
        p.n_tiles, p.tiles = int(self.d_tiles.shape[0]), self.d_tiles.data_ptr()
        if self.presort:
            p.n_tiles_coarse, p.tiles_coarse = int(self.d_tiles_coarse.shape[0]), self.d_tiles_coarse.data_ptr()
            if self.prune and self.d_boxes is not None:
                p.boxes = self.d_boxes.data_ptr()
        p.inv_band_scale = self.d_inv_scale.data_ptr()
        p.err_a, p.err_b = s.err_a, s.err_b
        p.grid, p.params = self.grid, self.params
        if want_pdf2d and self.d_pair_bin is not None:
            p.n_pairs, p.pair_len = len(s.pairs), int(s.pair_off[-1])
            p.pair_bin, p.pair_off = self.d_pair_bin.data_ptr(), self.d_pair_off.data_ptr()
        p.sm_count = self.sm_count
        p.sort_band = self.sort_band
        p.force_legacy = int(self.force_legacy or not self.split)
        p.max_range_entries = int(getattr(self, "max_range_entries", 0))
        return p

    def _workspace_for(self, plan, n):
        """one device buffer per engine, grown on demand: the fixed part of a batch of n stars
        plus ``scratch_budget_bytes`` for the lists and the per-range scratch"""
        torch = _torch()
        need = int(self.lib.bf_fit_workspace_bytes(C.byref(plan), n, 0)) + self.scratch_budget_bytes
        if self._workspace is None or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._workspace

    def stage_ms(self, reset=True):
        """device time per stage (ms, summed over the calls made with ``profile = True`` since the
        last reset) and call counts, as measured inside bf_fit_batch"""
        tot, cnt = dict(self._stage_tot), dict(self._stage_cnt)
        if reset:
            self._stage_tot, self._stage_cnt = {}, {}
        return tot, cnt

    def fit_batch(self, flux, threshold=-10.0, percentiles=(16.0, 50.0, 84.0), want_pdf2d=False,
                  sparse=None, rng=None, lnp_rng="device", seed=0, star_base=0, while_running=None):
        """Fit one batch (<= max_batch stars) with one call of ``bf_fit_batch``.  ``flux``:
        (n, nb) float64 NumPy array (host: copied inside the call) or CUDA tensor.  Returns a
        dict of device tensors, one row per star in input order.

        ``while_running``: called once after the C call has returned and before the host waits
        for the batch (the sparse pass is then still running on the device): host work to hide.

        ``sparse``: None, "all" or lnp_npts (fit.py:693-709).  ``lnp_rng="device"``: the subsample
        is drawn inside the call by a keyed permutation (same distribution as the reference,
        its own random stream).  ``lnp_rng="numpy"``: ranks from ``np.random.choice(n, size,
        replace=False)`` in catalog order, the call the reference makes, so a seeded run
        reproduces the reference's selection (O(|S|) host work per star)."""
        torch = _torch()
        s = self.setup
        if len(percentiles) != 3:
            raise ValueError("three percentiles per call")
        with torch.cuda.device(self.device):
            on_host = not (hasattr(flux, "is_cuda") and flux.is_cuda)
            if on_host:
                fl = np.ascontiguousarray(flux.numpy() if hasattr(flux, "numpy") else flux, dtype=np.float64)
                n, fptr = fl.shape[0], fl.ctypes.data
            else:
                fl = flux.to(self.device, dtype=torch.float64).contiguous()
                n, fptr = fl.shape[0], fl.data_ptr()
            if n > self.max_batch:
                raise ValueError("batch of %d stars exceeds max_batch=%d" % (n, self.max_batch))
            dev = self.device
            plan = self._plan(want_pdf2d)
            ws = self._workspace_for(plan, n)
            f8, i8 = torch.float64, torch.int64
            o = {
                "Pmax": torch.empty(n, dtype=f8, device=dev), "Pmax_j": torch.empty(n, dtype=i8, device=dev),
                "total_log_norm": torch.empty(n, dtype=f8, device=dev),
                "chi2min": torch.empty(n, dtype=f8, device=dev),
                "chi2min_j": torch.empty(n, dtype=i8, device=dev),
                "n_sparse": torch.empty(n, dtype=torch.int32, device=dev),
                "n_groups_selected": torch.empty(n, dtype=i8, device=dev),
                "exp_vals": torch.empty((n, s.n_q), dtype=f8, device=dev),
                "pdf1d": torch.empty((n, s.total_bins), dtype=f8, device=dev),
                "per_vals": torch.empty((n, s.n_q, 3), dtype=f8, device=dev),
            }
            oc = _cabi.FitOutputsC()
            oc.pmax, oc.pmax_j = o["Pmax"].data_ptr(), o["Pmax_j"].data_ptr()
            oc.total_log_norm, oc.chi2min = o["total_log_norm"].data_ptr(), o["chi2min"].data_ptr()
            oc.chi2min_j, oc.n_sparse = o["chi2min_j"].data_ptr(), o["n_sparse"].data_ptr()
            oc.n_groups_selected = o["n_groups_selected"].data_ptr()
            oc.exp_vals, oc.pdf1d, oc.per_vals = o["exp_vals"].data_ptr(), o["pdf1d"].data_ptr(), o["per_vals"].data_ptr()
            if plan.n_pairs:
                o["pdf2d"] = torch.empty((n, int(plan.pair_len)), dtype=f8, device=dev)
                oc.pdf2d = o["pdf2d"].data_ptr()
            mode, npts = 0, 0
            if sparse is not None and sparse != "all" and lnp_rng == "device":
                mode, npts = 2, int(sparse)
                o["sparse_off"] = torch.empty(n + 1, dtype=i8, device=dev)
                o["sparse_j"] = torch.empty(n * npts, dtype=i8, device=dev)
                o["sparse_lnp"] = torch.empty(n * npts, dtype=torch.float32, device=dev)
                o["sparse_chi2"] = torch.empty(n * npts, dtype=torch.float32, device=dev)
                oc.sparse_off, oc.sparse_j = o["sparse_off"].data_ptr(), o["sparse_j"].data_ptr()
                oc.sparse_lnp, oc.sparse_chi2 = o["sparse_lnp"].data_ptr(), o["sparse_chi2"].data_ptr()
                oc.sparse_capacity = n * npts
            state = _cabi.FitState()
            state.want_timing = int(self.profile)
            pct = (C.c_double * 3)(*[float(x) for x in percentiles])
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            _cabi.call("bf_fit_batch", C.byref(plan), n, C.c_void_p(fptr), int(on_host), float(threshold),
                       pct, mode, npts, int(seed) & 0xFFFFFFFFFFFFFFFF, int(star_base),
                       C.c_void_p(ws.data_ptr()), ws.numel(), C.byref(oc), C.byref(state), st,
                       C.c_void_p(self._aux_stream.cuda_stream))
            self.launches += int(state.launches)
            self.last_state = (plan, state)
            if self.profile:
                for name, v in zip(_cabi.STAGES, state.stage_ms):
                    self._stage_tot[name] = self._stage_tot.get(name, 0.0) + float(v)
                    self._stage_cnt[name] = self._stage_cnt.get(name, 0) + 1
            if while_running is not None:
                while_running()
            if mode == 2:
                nrec = int(o["sparse_off"][-1].item())
                for name in ("sparse_j", "sparse_lnp", "sparse_chi2"):
                    o[name] = o[name][:nrec]
            elif sparse is not None:
                self._host_ranked_records(o, sparse, rng, plan, state, n, st)
            return o

    def _host_ranked_records(self, o, sparse, rng, plan, state, n, st):
        """sparse records for "all" (every member) or for ranks drawn on the host with NumPy's
        stream (fit.py:696), after a bf_fit_batch call without records"""
        torch = _torch()
        dev = self.device
        ns_host = (o["n_sparse"].to(torch.int64) & 0xFFFFFFFF).cpu().numpy()
        ranks = None
        if sparse == "all":
            cnt = ns_host
        else:
            npts = int(sparse)
            choose = rng.choice if rng is not None else np.random.choice
            parts = [np.asarray(choose(int(m), size=npts, replace=False)) if npts < m else np.arange(m)
                     for m in ns_host]
            cnt = np.minimum(ns_host, npts)
            if len(parts) and int(cnt.sum()):
                ranks = torch.from_numpy(np.concatenate(parts).astype(np.int32)).to(dev)
        rec_off = np.zeros(n + 1, dtype=np.int64)
        rec_off[1:] = np.cumsum(cnt)
        nrec = int(rec_off[-1])
        o["sparse_off"] = torch.from_numpy(rec_off).to(dev)
        o["sparse_j"] = torch.empty(max(nrec, 1), dtype=torch.int64, device=dev)
        o["sparse_lnp"] = torch.empty(max(nrec, 1), dtype=torch.float32, device=dev)
        o["sparse_chi2"] = torch.empty(max(nrec, 1), dtype=torch.float32, device=dev)
        if nrec:
            _cabi.call("bf_fit_sparse_records", C.byref(plan), C.byref(state), n, _ptr(o["sparse_off"]),
                       _ptr(ranks), _ptr(o["sparse_j"]), _ptr(o["sparse_lnp"]), _ptr(o["sparse_chi2"]), st)
            self.launches += 1
        for name in ("sparse_j", "sparse_lnp", "sparse_chi2"):
            o[name] = o[name][:nrec]

    # ------------------------------------------------------------------------------
    def fit(self, fluxes, threshold=-10.0, percentiles=(16.0, 50.0, 84.0), want_pdf2d=False,
            sparse=None, batch=None, rng=None, lnp_rng="device", seed=0, reuse_buffers=False,
            result=None, row_offset=0, on_batch=None, tail_batch=0):
        """Fit all stars of ``fluxes`` (ns, nb) in batches; returns a host ``FitResult``.

        Results are copied device->host asynchronously into page-locked arrays while the next
        batch computes.  With ``reuse_buffers=True`` the engine keeps those arrays and hands
        the same ones back on the next call of the same shape (they are overwritten then).
        ``result`` / ``row_offset``: write into rows ``row_offset ...`` of an existing
        ``FitResult`` (a resumed run); ``on_batch(lo, hi)`` is called after the rows [lo, hi) of
        a batch have arrived on the host (checkpoints).  ``tail_batch``: when there are several
        batches, end with one of at most this many stars -- the copy of the last batch's results
        is the only one no computation hides, so a short last batch shortens the call."""
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
        res = result
        if res is None:
            key = (ns, pair_len, cap)
            res = self._result_cache.get(key) if reuse_buffers else None
            if res is None:
                res = FitResult(ns, s.n_q, s.total_bins, pair_len, cap)
                if reuse_buffers:
                    self._result_cache = {key: res}
        t = res._t
        sp_parts = []
        sp_pos = res.sparse_filled if result is not None else 0
        if sparse is not None:
            if result is None or res.sparse_off is None:
                res.sparse_off = np.zeros(res.ns + 1, dtype=np.int64)
            if "_sp_j" in t:  # records land in the preallocated arrays as the batches finish
                res.sparse_j, res.sparse_lnp, res.sparse_chi2 = res._sp_j, res._sp_lnp, res._sp_chi2
        starts = list(range(0, ns, bs))
        if tail_batch and len(starts) > 1 and ns - starts[-1] > 2 * int(tail_batch):
            starts.append(ns - int(tail_batch))
        ends = starts[1:] + [ns]
        n_batches = len(starts)
        # record offsets of every batch, staged in page-locked memory: they are small and go first
        # on the copy stream, so that nothing on the host waits behind a batch's 1D PDFs
        off_stage = torch.empty((n_batches, bs + 1), dtype=torch.int64, pin_memory=True) if sparse is not None else None
        settled = [0]
        spans = []

        def settle(upto):  # offsets of the batches [settled, upto), whose copies have arrived
            off = res.sparse_off
            for i in range(settled[0], upto):
                lo, hi = spans[i]
                off[lo + 1:hi + 1] = off[lo] + off_stage[i, 1:hi - lo + 1].numpy()
                off[hi + 1:] = off[hi]
            settled[0] = max(settled[0], upto)

        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream(self.device)
            copy_stream = self._copy_stream
            pending = None
            def previous_batch_done():
                # the previous batch's rows have had a C call's worth of time to arrive; the host
                # work on them runs while this batch's sparse pass does
                nonlocal pending
                if on_batch is not None and pending is not None:
                    pending[0].synchronize()
                    if sparse is not None:
                        settle(pending[3] + 1)
                    on_batch(pending[1], pending[2])
                    pending = None

            for i, (a, b) in enumerate(zip(starts, ends)):
                o = self.fit_batch(fluxes[a:b], threshold, percentiles, want_pdf2d, sparse, rng,
                                   lnp_rng, seed, row_offset + a, while_running=previous_batch_done)
                done = torch.cuda.Event()
                done.record(main)
                copy_stream.wait_event(done)
                lo, hi = row_offset + a, row_offset + b
                spans.append((lo, hi))
                with torch.cuda.stream(copy_stream):
                    if sparse is not None:
                        o["sparse_off"].record_stream(copy_stream)
                        off_stage[i, :hi - lo + 1].copy_(o["sparse_off"], non_blocking=True)
                    names = ["Pmax", "Pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse",
                             "n_groups_selected", "exp_vals", "per_vals", "pdf1d"]
                    if pair_len:
                        names.append("pdf2d")
                    for name in names:
                        o[name].record_stream(copy_stream)
                        t[name][lo:hi].copy_(o[name], non_blocking=True)
                    if sparse is not None:
                        nrec = int(o["sparse_j"].shape[0])
                        if "_sp_j" in t and sp_pos + nrec <= t["_sp_j"].shape[0]:
                            for name, src in (("_sp_j", "sparse_j"), ("_sp_lnp", "sparse_lnp"),
                                              ("_sp_chi2", "sparse_chi2")):
                                o[src].record_stream(copy_stream)
                                t[name][sp_pos:sp_pos + nrec].copy_(o[src], non_blocking=True)
                            sp_pos += nrec
                        else:
                            sp_parts.append([o["sparse_j"].cpu(), o["sparse_lnp"].cpu(),
                                             o["sparse_chi2"].cpu()])
                    arrived = torch.cuda.Event()
                    arrived.record(copy_stream)
                del o  # (record_stream keeps the device buffers alive until their copies are done)
                if on_batch is not None:
                    previous_batch_done()  # (normally done already, inside fit_batch)
                    pending = (arrived, lo, hi, i)
            copy_stream.synchronize()
            main.synchronize()
            if sparse is not None:
                settle(n_batches)
            if on_batch is not None and pending is not None:
                on_batch(pending[1], pending[2])
        if sparse is not None:
            if not sp_parts:
                n = sp_pos
                res.sparse_filled = n
                res.sparse_j, res.sparse_lnp, res.sparse_chi2 = res._sp_j[:n], res._sp_lnp[:n], res._sp_chi2[:n]
            else:
                res.sparse_j = np.concatenate([p[0].numpy() for p in sp_parts])
                res.sparse_lnp = np.concatenate([p[1].numpy() for p in sp_parts])
                res.sparse_chi2 = np.concatenate([p[2].numpy() for p in sp_parts])
        return res
