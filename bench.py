#!/usr/bin/env python
"""
Benchmark of the BEAST fitting hot path on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2]

A *step* is one pass of the hot path over the whole synthetic catalog of the named
configuration (C2 = PHAT-like 6-band grid of 4e6 models, 1e5 stars, stats + pdf1d +
sparse lnP).  ``value`` = star x model lnP evaluations per second with the fluxes already
resident in HBM; ``e2e`` = the same through the public engine API with host fluxes in and
all results back on the host, every step.  For N > 1 every rank fits its own 1e5 stars
against a replicated grid (stars shard, no data-path collective): weak scaling.

``--impl reference`` times the CPU restatement of the reference's star loop
(oracle/fit_oracle.py, faithful mode) on the box's host cores, one process per core over
disjoint star slices -- the reference's own parallel model (tools/run/helper_functions.py:55-58).
"""
import argparse
import contextlib
import json
import os
import sys
import threading
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (axes name, stars per GPU, description)
    "C1": ("C1", 1000, "synthetic PHAT 6-band SEDGrid 1e5 models, 1,000 stars"),
    "C2": ("C2", 100000, "PHAT 6-band 4e6-model grid, 1e5 stars, stats+pdf1d+sparse lnP"),
    "C3": ("C2", 1000000, "PHAT 6-band 4e6-model grid, 1e6 stars sharded over the GPUs (grid replicated)"),
    "C4": ("C4", 200000, "8-band UV-NIR grid 1e7 models, 2e5 stars"),
    "C5": ("C5", 1000000, "2e7-model grid as 8 subgrids, NCCL merge of lnP/pdf1d/best-fit, 1e6 stars"),
}
THRESHOLD = -10.0  # summary_table_memory / run_fitting default (fit.py:924, run_fitting.py:303)
MAX_NBINS = 200    # summary_table_memory default (fit.py:928)
LNP_NPTS = 500     # run_fitting.py:305


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--stars", type=int, default=None, help="override stars per GPU")
    ap.add_argument("--max-nbins", type=int, default=MAX_NBINS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pdf2d", action="store_true", help="also build 2D PDFs for Av, M_ini, logT (config C4)")
    ap.add_argument("--cpu-stars-per-core", type=int, default=2)
    ap.add_argument("--files", action="store_true",
                    help="also time one plugin call that writes stats/pdf1d/lnp files to tmpfs")
    ap.add_argument("--sparse-pass", default="split", choices=["split", "legacy", "both"],
                    help="kernels of the sparse pass (FitEngine sparse_pass); measurements")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU during the timed region (NVML)"""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
               0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
               0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz, self.ok = None, False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        while self.ok and not self.stop_flag:
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                bits = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                if getattr(self, "recording", True):
                    self.samples.append(mhz)
                    for b, name in self.REASONS.items():
                        if bits & b and name != "gpu_idle":
                            self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml_unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            return json.load(f), "MEASURED_PEAKS.json"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def build_inputs(cfg, stars, seed):
    from beast_b200 import synth

    axes = CONFIGS[cfg][0]
    if axes == "C4":
        g = synth.make_sedgrid(axes, synth.UVNIR_FILTERS, synth.UVNIR_LAMB, seed=1234)
    else:
        g = synth.make_sedgrid(axes, seed=1234)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, stars, seed=seed, with_coords=False)
    return g, nz, obs


# ---------------------------------------------------------------------------------------
_CPU = {}
_OUT = sys.stdout


def _cpu_worker(args):
    lo, hi = args
    from oracle import fit_oracle as fo

    st, fluxes = _CPU["setup"], _CPU["flux"]
    t = time.perf_counter()
    for e in range(lo, hi):
        fo.fit_star(st, fluxes[e], THRESHOLD, fast=False)
    return time.perf_counter() - t


def _ref_worker(args):
    lo, hi, block = args
    from oracle import ref_shim

    return ref_shim.time_reference_star_loop(_CPU["flux"][lo:hi], _CPU["g"], _CPU["nz"], _CPU["keys"],
                                             block, THRESHOLD, _CPU["max_nbins"])


def cpu_reference_itself(g, nz, obs, max_nbins, stars_per_core, steps, warmup):
    """time the star loop of the UNMODIFIED reference (its files under oracle/_ref, or the tree
    at /root/reference) on all host cores: one process per core runs ``Q_all_memory`` once over
    its own stars; the loop is timed block by block from inside (oracle/ref_shim.py), so the
    setup of the bins is outside the timed region, as for the GPU arm."""
    import multiprocessing as mp

    from oracle import fit_oracle as fo
    from oracle import ref_shim

    if not ref_shim.reference_available():
        raise RuntimeError("no reference files (oracle/_ref is built by oracle/make_ref.py)")
    cores = len(os.sched_getaffinity(0))
    per = (warmup + steps) * stars_per_core
    if len(obs) < cores * per:
        raise RuntimeError("catalog too small for the CPU sample")
    _CPU.update(g=g, nz=nz, keys=fo.default_keys(g), flux=obs.fluxes, max_nbins=max_nbins)
    nm = int(np.count_nonzero(g["weight"] > 0))
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        blocks = pool.map(_ref_worker, [(c * per, (c + 1) * per, stars_per_core) for c in range(cores)])
    rates = []
    for it in range(warmup, warmup + steps):
        dt = max(float(b[it]) for b in blocks)  # the step ends when the slowest core is done
        rates.append((cores * stars_per_core * nm / dt, dt))
    sample = ("%d stars (%d per core x %d cores) of the same catalog against the full %d-model grid per step, "
              "fitted by the reference's own Q_all_memory (star loop timed from inside, setup excluded)"
              % (cores * stars_per_core, stars_per_core, cores, nm))
    return rates, cores, sample, "reference"


def cpu_reference_run(g, nz, obs, max_nbins, stars_per_core, steps, warmup):
    """time the oracle's faithful star loop on all host cores; returns (evals/s per step list,
    cores, sample description).  Setup (bins, index lists) is outside the timed region."""
    import multiprocessing as mp

    from oracle import fit_oracle as fo

    try:  # the reference itself where its files are at hand, else the port
        return cpu_reference_itself(g, nz, obs, max_nbins, stars_per_core, steps, warmup)
    except Exception as exc:
        sys.stderr.write("bench: reference files not usable (%r), timing the port\n" % (exc,))
    cores = len(os.sched_getaffinity(0))
    keys = fo.default_keys(g)
    st = fo.OracleSetup(g, nz, keys, max_nbins=max_nbins)
    for b in st.bins:
        b.index_lists()
    _CPU["setup"], _CPU["flux"] = st, obs.fluxes
    nm = len(st.g0_indxs)
    rates = []
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        for it in range(warmup + steps):
            base = (it * cores * stars_per_core) % max(1, len(obs) - cores * stars_per_core)
            jobs = [(base + c * stars_per_core, base + (c + 1) * stars_per_core) for c in range(cores)]
            t = time.perf_counter()
            pool.map(_cpu_worker, jobs)
            dt = time.perf_counter() - t
            if it >= warmup:
                rates.append((cores * stars_per_core * nm / dt, dt))
    sample = "%d stars (%d per core x %d cores) of the same catalog against the full %d-model grid per step" % (
        cores * stars_per_core, stars_per_core, cores, nm)
    return rates, cores, sample, "port"


# ---------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    if rank != 0:
        return
    stars = max(64, len(os.sched_getaffinity(0)) * args.cpu_stars_per_core * (args.steps + args.warmup))
    g, nz, obs = build_inputs(args.config, stars, seed=4321)
    rates, cores, sample, kind = cpu_reference_run(g, nz, obs, args.max_nbins, args.cpu_stars_per_core,
                                                   args.steps, args.warmup)
    nm = int(np.count_nonzero(g["weight"] > 0))
    val = float(np.mean([r for r, _ in rates]))
    ms = float(np.mean([dt for _, dt in rates]) * 1e3)
    line = {
        "impl": "reference", "metric": "star-model lnP evaluations per second", "value": val,
        "unit": "star*model/s", "stars_per_s": val / nm, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, nm, stars),
        "cpu_baseline": {"value": val, "unit": "star*model/s", "cores": cores, "kind": kind,
                         "sample": sample},
        "e2e": {"value": val, "unit": "star*model/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()


def workload_config(args, nm, stars):
    return {"workload": CONFIGS[args.config][2], "config": args.config, "models": nm,
            "bands": 8 if args.config == "C4" else 6, "stars_per_gpu": stars,
            "threshold": THRESHOLD, "max_nbins": args.max_nbins, "lnp_npts": LNP_NPTS,
            "outputs": "stats+pdf1d+sparse lnP" + ("+pdf2d" if args.pdf2d else ""), "parallelism": "stars sharded, grid replicated",
            "l2": "inputs larger than L2 (grid tiles + exact arrays + parameter tables > 126 MB)"}


def run_ours(args, rank, world, local_rank):
    import torch

    from beast_b200 import fit as bfit
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from beast_b200.engine import bind_host_to_gpu

    # one process per GPU, on the cores (and memory) of the socket that GPU hangs on
    host_cores = bind_host_to_gpu(local_rank) if os.environ.get("BF_NO_BIND") is None else None
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)

    # C3 is C2's grid with 1e6 stars split over the ranks (strong scaling); the others give
    # every rank its own catalog (weak scaling)
    strong = args.config == "C3"
    total_stars = args.stars or CONFIGS[args.config][1]
    stars = -(-total_stars // world) if strong else total_stars
    g, nz, obs = build_inputs(args.config, stars, seed=4321 + rank)
    keys = [k for k in g.keys() if k not in
            "osl keep weight grid_weight prior_weight fullgrid_idx stage specgrid_indx phase eep".split()]
    t0 = time.perf_counter()
    p2d = ["Av", "M_ini", "logT"] if args.pdf2d else None
    setup = FitSetup(g, nz, keys, max_nbins=args.max_nbins, pdf2d_param_list=p2d, make_pdf2d=args.pdf2d)
    eng = FitEngine(setup, device=dev, gmin_budget_bytes=24 << 30, sparse_pass=args.sparse_pass)
    setup_s = time.perf_counter() - t0
    nm = setup.n_models
    fluxes = np.ascontiguousarray(obs.fluxes)
    bs = eng.max_batch
    if args.pdf2d:  # 2D PDFs are dense per star: same bound as FitEngine.fit
        bs = int(max(1, min(bs, (16 << 30) // (int(setup.pair_off[-1]) * 8))))
    batches = [(a, min(stars, a + bs)) for a in range(0, stars, bs)]
    d_flux = [torch.from_numpy(fluxes[a:b]).to(dev) for a, b in batches]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_resident():
        for (a, b), df in zip(batches, d_flux):
            eng.fit_batch(df, THRESHOLD, sparse=LNP_NPTS, lnp_rng="device", seed=1, star_base=a,
                          want_pdf2d=args.pdf2d)

    # ---- device-resident throughput: one bf_fit_batch call per batch of stars ------------
    sampler = ClockSampler(local_rank)  # started before the warm-up: the first NVML queries stall
    sampler.recording = False           # CUDA API calls for a while on some drivers
    sampler.start()
    for _ in range(args.warmup):
        step_resident()
    barrier()
    eng.launches = 0
    sampler.recording = True
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step_resident()
    ev1.record()
    barrier()
    sampler.stop_flag = True
    ms_total = ev0.elapsed_time(ev1)
    launches = eng.launches
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = world * stars * nm / (ms_step * 1e-3)

    # ---- one more step with bf_fit_batch timing its stages (CUDA events on the launching stream,
    # a synchronisation after every stage: outside the timed region) -------------------------
    eng.profile = True
    eng.stage_ms()
    step_resident()
    torch.cuda.synchronize(dev)
    stage_tot, stage_cnt = eng.stage_ms()
    eng.profile = False
    n_list = 0
    dense_evaluated = 0
    for (a, b), df in zip(batches, d_flux):  # listed groups of the catalog (for the rooflines)
        eng.fit_batch(df, THRESHOLD)
        n_list += int(eng.last_state[1].n_list)
        dense_evaluated += int(eng.last_state[1].dense_evaluated)

    # ---- end to end through the reference-facing call: Q_all_memory with host objects in and
    # the reference's result arrays back on the host, every step ------------------------------
    e2e_stars = stars
    if args.pdf2d:  # 2D PDFs of a whole catalog do not fit host memory (1 MB per star)
        e2e_stars = min(stars, 16384)
    # the result arrays are page-locked host memory: at most ~16 GB of 1D PDFs per call
    e2e_stars = int(min(e2e_stars, max(1024, (16 << 30) // (setup.total_bins * 8))))
    e2e_obs = obs if e2e_stars == stars else obs.subset(slice(0, e2e_stars))
    eng.reuse_buffers = True  # page-locked result arrays are kept across calls (not re-pinned)
    qkw = dict(p=[16.0, 50.0, 84.0], max_nbins=args.max_nbins, lnp_npts=LNP_NPTS, threshold=THRESHOLD,
               engine=eng, keep_lnp=True, pdf2d_param_list=p2d)

    def step_e2e(files=None):
        kw = dict(qkw)
        if files:
            kw.update(stats_outname=files + "_stats.fits", pdf1d_outname=files + "_pdf1d.fits",
                      lnp_outname=files + "_lnp.hd5", keep_lnp=False)
        if args.pdf2d:
            kw["pdf2d_outname"] = (files + "_pdf2d.fits") if files else os.devnull
        with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
            return bfit.Q_all_memory({}, e2e_obs, g, nz, list(keys), **kw)

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = step_e2e()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    h2d = e2e_obs.fluxes.nbytes
    res = eng._result_cache[next(iter(eng._result_cache))]
    d2h = sum(getattr(res, n)[:e2e_stars].nbytes for n in ("Pmax", "Pmax_j", "total_log_norm", "chi2min",
                                                          "chi2min_j", "n_sparse", "n_groups_selected",
                                                          "exp_vals", "per_vals", "pdf1d"))
    d2h += sum(getattr(res, n).nbytes for n in ("sparse_j", "sparse_lnp", "sparse_chi2")) + 8 * e2e_stars
    if args.pdf2d:
        d2h += res.pdf2d[:e2e_stars].nbytes
    mean_sparse = float(np.mean(res.n_sparse[:e2e_stars]))
    mean_groups = float(np.mean(res.n_groups_selected[:e2e_stars]))
    members = float(np.sum(res.n_sparse[:e2e_stars].astype(np.float64))) * (stars / e2e_stars)
    files_s = None
    if rank == 0 and args.files and not args.pdf2d:
        import tempfile

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as td:
                t0 = time.perf_counter()
                step_e2e(os.path.join(td, "bench"))
                files_s = time.perf_counter() - t0
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- rooflines ------------------------------------------------------------------------------
    peaks, peak_src = measured_peaks()
    nb = setup.nb
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    sm_max = float(peaks.get("sm_max_mhz", 1965.0))
    peak_tf = eng.sm_count * 128 * 2 * sm_max * 1e6 / 1e12
    flop_per_eval = 4 * nb + 2  # 2*nb+1 FMA per star x model
    nbat = len(batches)

    def per_launch(name):
        return stage_tot.get(name, 0.0) / max(nbat, 1)

    dense_ms = per_launch("dense")
    ld = -(-int(stars / nbat) // 1024) * 1024
    # the pruned sweep reports the (star, model) pairs it evaluated (padding included); the
    # roofline counts those, not the pairs it could rule out
    evals_all = float(nbat) * ld * setup.ld_model
    evals_per_launch = dense_evaluated / nbat
    dense_tf = flop_per_eval * evals_per_launch / (dense_ms * 1e-3) / 1e12 if dense_ms else None
    dense_traffic = (ld // 1024) * setup.tiles.nbytes + setup.n_groups * ld * 4
    peak_note = "%d SMs x 128 FP32 lanes x 2 x sm_max_mhz (%s); MEASURED_PEAKS.json has no FP32 entry" % (
        eng.sm_count, peak_src)
    r_dense = {"bound": "fp32", "kernel": "dense_chi2_kernel", "achieved": dense_tf, "peak": peak_tf,
               "unit": "TFLOP/s", "frac": (dense_tf / peak_tf) if dense_tf else None,
               "traffic": int(dense_traffic), "ms_per_step": stage_tot.get("dense", 0.0),
               "work": "%d flop per evaluated star x model (2 nb + 1 FMA); the sweep evaluated %.3g of %.3g "
                       "pairs (the others ruled out per group by a lower bound, 5 nb + 2 flop per star x group, "
                       "not counted)" % (flop_per_eval, dense_evaluated, evals_all),
               "evaluated_fraction": dense_evaluated / evals_all if evals_all else None,
               "peak_source": peak_note,
               "traffic_note": "bytes per launch = star blocks x tile bytes + group-minimum map; equals ncu "
                               "dram__bytes of the kernel (profiles/r01_dense_chi2_ncu.md)"}
    # exact weights: every listed model's FP64 rows (2 nb + 2 doubles) are needed once, every
    # member's weight is produced once (SURVEY.md 8d / VERDICT r01: 112 B x evaluated + 8 B x members)
    w_ms = stage_tot.get("exact_weights", 0.0)
    w_bytes = (2 * nb + 2) * 8.0 * 64 * n_list + 8.0 * members
    r_weights = {"bound": "hbm", "kernel": "exact_weights_grouped_kernel" if eng.split else "exact_weights_kernel",
                 "achieved": w_bytes / (w_ms * 1e-3) / 1e9 if w_ms else None, "peak": hbm, "unit": "GB/s",
                 "frac": (w_bytes / (w_ms * 1e-3) / 1e9 / hbm) if w_ms else None, "traffic": None,
                 "ms_per_step": w_ms,
                 "work": "%d B of model rows per listed model + 8 B per member weight; the kernel writes 64 weights "
                         "(+ run sums, expectation partials) per listed group, 0.7 KB, which is what its DRAM "
                         "traffic consists of (profiles/r02_exact_weights_grouped_ncu.md)" % ((2 * nb + 2) * 8)}
    # histograms: every member's weight is read once and added to one bin of each parameter
    h_ms = stage_tot.get("histograms", 0.0)
    h_bytes = 8.0 * members + stars * setup.total_bins * 8.0
    upd = members * setup.n_q
    r_hist = {"bound": "hbm", "kernel": "split_hist_kernel" if eng.split else "hist_window_kernel / hist_kernel",
              "achieved": h_bytes / (h_ms * 1e-3) / 1e9 if h_ms else None, "peak": hbm, "unit": "GB/s",
              "frac": (h_bytes / (h_ms * 1e-3) / 1e9 / hbm) if h_ms else None, "traffic": None,
              "ms_per_step": h_ms,
              "work": "8 B per member weight in + one pdf row per star out; the kernel is bound by "
                      "shared-memory updates, not by HBM: %.3g histogram updates per step = %.3g per second"
                      % (upd, upd / (h_ms * 1e-3) if h_ms else 0.0),
              "updates_per_s": upd / (h_ms * 1e-3) if h_ms else None}
    kernels = [r_dense, r_weights, r_hist]
    dominant = max(kernels, key=lambda r: r["ms_per_step"] or 0.0)
    roofline = dict(dominant)
    roofline["kernels"] = kernels
    roofline["hbm_gbs_peak"] = hbm
    stage_sum = sum(stage_tot.values()) or 1.0
    roofline["stage_ms_per_step"] = dict(stage_tot)
    roofline["stage_share_of_step"] = {k: v / stage_sum for k, v in stage_tot.items()}
    roofline["stage_note"] = ("stage times from one extra step in which bf_fit_batch synchronises after every "
                              "stage; their sum is %.1f ms against %.1f ms for the timed step" % (stage_sum, ms_step))
    # the metric BASELINE.json states: fraction of the FP32 pipe peak for the whole fit, i.e. the
    # reference's work (every star against every model) per second over the peak -- with the pruned
    # sweep this is a rate of results, not a pipe utilisation (that is r_dense["frac"])
    roofline["frac_of_fp32_peak"] = flop_per_eval * stars * nm / (ms_step * 1e-3) / 1e12 / peak_tf

    # ---- the step before the fit: trim_models flags pass, HBM-bound (fitting/trim_grid.py) -----
    trim_info = None
    try:
        import ctypes as C

        from beast_b200 import _cabi

        up = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)  # noqa: E731
        t_seds, t_bias, t_unc, t_cmp = up(g.seds), up(nz["bias"]), up(nz["error"]), up(nz["completeness"])
        lo, hi = up(fluxes.min(axis=0)), up(fluxes.max(axis=0))
        pat = torch.empty(len(g), dtype=torch.int32, device=dev)
        hist = torch.zeros(1 << (setup.nb + 1), dtype=torch.int32, device=dev)
        P = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        ta, tb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = []
        for it in range(5):
            ta.record()
            _cabi.call("bf_trim_flags", setup.nb, len(g), P(t_seds), P(t_bias), P(t_unc), P(t_cmp), 3.0,
                       P(lo), P(hi), P(pat), P(hist), st)
            tb.record()
            torch.cuda.synchronize(dev)
            ms.append(ta.elapsed_time(tb))
        nbytes = 4 * t_seds.numel() * 8 + 4 * len(g)
        best = min(ms[1:])
        trim_info = {"kernel": "trim_flags_kernel", "models": len(g), "ms": best,
                     "achieved_gbs": nbytes / best / 1e6, "peak_gbs": hbm,
                     "frac": nbytes / best / 1e6 / hbm, "bound": "hbm"}
        del t_seds, t_bias, t_unc, t_cmp
    except Exception as exc:  # the fit numbers above stand on their own
        trim_info = {"error": repr(exc)}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        rates, cores, sample, kind = cpu_reference_run(g, nz, obs, args.max_nbins, args.cpu_stars_per_core, 1, 1)
        cpu = {"value": rates[0][0], "unit": "star*model/s", "cores": cores, "kind": kind,
               "sample": sample, "seconds": rates[0][1]}

    cfg = workload_config(args, nm, stars)
    if strong:
        cfg["parallelism"] = "1e6 stars split over the GPUs, grid replicated"
        cfg["total_stars"] = stars * world
    line = {
        "metric": "star-model lnP evaluations per second", "value": value, "unit": "star*model/s",
        "stars_per_s": value / nm, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32 dense sweep + f64 exact pass", "data": "synthetic",
        "config": cfg,
        "clocks": sampler.summary(),
        "e2e": {"value": world * e2e_stars * nm / e2e_s, "unit": "star*model/s",
                "stars_per_s": world * e2e_stars / e2e_s, "seconds_per_step": e2e_s,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "stars": e2e_stars,
                "call": "beast_b200.fit.Q_all_memory(prev_result, obs, sedgrid, obsmodel, keys, p=[16,50,84], "
                        "max_nbins=%d, lnp_npts=%d, threshold=%g, engine=<resident grid>, keep_lnp=True): host "
                        "objects in, the reference's result arrays (stats, pdf1d, sparse lnP) back on the host, "
                        "no files" % (args.max_nbins, LNP_NPTS, THRESHOLD),
                "with_files_on_tmpfs_seconds": files_s,
                "host_cores_rank0": ("%d cores next to the GPU (NVML affinity)" % len(host_cores)) if host_cores
                else "unbound"},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "cpu_baseline": cpu,
        "trim": trim_info,
        "setup_seconds": setup_s,
        "mean_sparse_set": mean_sparse,
        "mean_groups_selected": mean_groups,
    }
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()
    if dist is not None:
        dist.destroy_process_group()


def run_c5(args, rank, world, local_rank):
    """Config C5: the 2e7-model grid as 8 row slices (subgrids), every star fitted against every
    slice (do_not_normalize, shared grid_info_dict) and the per-star partial results merged per
    batch on the devices -- docs/subgrid_parallelism.rst:77-143 with the files replaced by an
    all_gather of three scalars per (star, subgrid) and a reduce_scatter by star of the weighted
    pdf1d / expectation sums (beast_b200.subgrid.SubgridFit).  N ranks hold 8/N slices each."""
    import torch
    import torch.distributed as dist

    from beast_b200 import subgrid, synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import SKIP_KEYS, FitSetup

    n_sub = 8
    if n_sub % world:
        raise SystemExit("C5 needs a number of GPUs that divides 8")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD if world > 1 else None
    L = n_sub // world
    axes = synth.CONFIG_AXES["C5"]
    n_total = int(np.prod(axes))
    slices = subgrid.uniform_slices(n_total, n_sub)
    t0 = time.perf_counter()
    mine = []
    for l in range(L):
        sl = slices[rank * L + l]
        g = synth.make_sedgrid("C5", seed=1234, rows=(sl.start, sl.stop))
        mine.append((g, synth.make_noisemodel(g)))
    infos = [subgrid.subgrid_info(g, nz) for g, nz in mine]
    for i in infos:  # only the number of unique values matters beyond the bin cap
        for q in i:
            if len(i[q]["unique"]) > 1000:
                i[q]["unique"] = i[q]["unique"][:1001]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, infos)
        infos = [x for part in gathered for x in part]
    info = subgrid.reduce_grid_info(infos)
    keys = [k for k in mine[0][0].keys() if k not in SKIP_KEYS]
    engines = []
    for g, nz in mine:
        st = FitSetup(g, nz, list(keys), max_nbins=args.max_nbins, grid_info_dict=info, do_not_normalize=True)
        engines.append(FitEngine(st, device=dev, gmin_budget_bytes=8 << 30, scratch_budget_bytes=8 << 30,
                                 sparse_pass=args.sparse_pass))
    setup_s = time.perf_counter() - t0
    stars = args.stars or 1000000
    # the same catalog on every rank: an equal share of the stars is drawn from every slice
    # (model index in proportion to the prior weight, observations.py:402-422), then interleaved
    per_sub = -(-stars // n_sub)
    parts = [synth.make_observations(g, nz, per_sub, seed=4321 + rank * L + l, with_coords=False).fluxes
             for l, (g, nz) in enumerate(mine)]
    loc = torch.from_numpy(np.ascontiguousarray(np.stack(parts))).to(dev)  # [L, per_sub, nb]
    if world > 1:
        allf = torch.empty((world,) + tuple(loc.shape), dtype=loc.dtype, device=dev)
        dist.all_gather_into_tensor(allf, loc)
        loc = allf.reshape(n_sub, per_sub, -1)
    fl = loc.permute(1, 0, 2).reshape(n_sub * per_sub, -1)[:stars].contiguous()
    fluxes = fl.cpu().numpy()
    bs = min(e.max_batch for e in engines)
    batches = [(a, min(stars, a + bs)) for a in range(0, stars, bs)]
    sf = subgrid.SubgridFit(engines, subgrid.CudaMergeMath(engines[0].setup), group, rank, world)
    n_models = sum(e.setup.n_models for e in engines)
    if world > 1:
        t = torch.tensor([n_models], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        n_models = int(t.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step(host_out=None, upto=None):
        pos = 0
        for a, b in batches:
            if upto is not None and a >= upto:
                break
            m = sf.fit_batch(fl[a:b] if host_out is None else fluxes[a:b], THRESHOLD)
            if host_out is not None:  # this rank's share of the batch's merged rows
                lo, hi = m["rows"]
                for name in ("pdf1d", "per_vals", "exp_vals", "best_vals", "Pmax", "total_log_norm", "chi2min"):
                    host_out[name][pos:pos + hi - lo].copy_(m[name], non_blocking=True)
                pos += hi - lo

    sampler = ClockSampler(local_rank)
    sampler.recording = False
    sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    for e in engines:
        e.launches = 0
    sf.collective_bytes = 0
    sampler.recording = True
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    sampler.stop_flag = True
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    launches = sum(e.launches for e in engines)
    coll_bytes = sf.collective_bytes / max(args.steps, 1)
    # one more step with the collectives and the stages of bf_fit_batch timed on the device
    # (synchronising: outside the timed region)
    sf.time_collectives, sf.collective_ms = True, 0.0
    for e in engines:
        e.profile = True
        e.stage_ms()
    step()
    barrier()
    coll_ms = sf.collective_ms
    sf.time_collectives = False
    stage_tot = {}
    for e in engines:
        tot, _ = e.stage_ms()
        e.profile = False
        for k, v in tot.items():
            stage_tot[k] = stage_tot.get(k, 0.0) + v
    # end to end: host fluxes in, this rank's share of the merged rows back on the host
    st0 = engines[0].setup
    # (page-locked result arrays for this rank's share: at most ~16 GB, i.e. whole batches of stars)
    cap_rows = max(1, (16 << 30) // (st0.total_bins * 8))
    e2e_batches = max(1, min(len(batches), cap_rows * world // bs))
    e2e_stars = min(stars, batches[e2e_batches - 1][1])
    rows = sum(-(-(b - a) // world) for a, b in batches[:e2e_batches])
    host = {"pdf1d": torch.empty((rows, st0.total_bins), dtype=torch.float64, pin_memory=True),
            "per_vals": torch.empty((rows, st0.n_q, 3), dtype=torch.float64, pin_memory=True)}
    for name in ("exp_vals", "best_vals"):
        host[name] = torch.empty((rows, st0.n_q), dtype=torch.float64, pin_memory=True)
    for name in ("Pmax", "total_log_norm", "chi2min"):
        host[name] = torch.empty(rows, dtype=torch.float64, pin_memory=True)
    step(host, e2e_stars)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(host, e2e_stars)
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.destroy_process_group()
    e2e_s = float(t.item())
    if rank != 0:
        return
    peaks, peak_src = measured_peaks()
    nb = st0.nb
    peak_tf = engines[0].sm_count * 128 * 2 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12
    value = stars * n_models / (ms_step * 1e-3)
    share = -(-stars // world)
    line = {
        "metric": "star-model lnP evaluations per second", "value": value, "unit": "star*model/s",
        "stars_per_s": stars / (ms_step * 1e-3), "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 dense sweep + f64 exact pass", "data": "synthetic",
        "config": {"workload": "2e7-model grid as 8 subgrids, NCCL merge of lnP/pdf1d/best-fit, 1e6 stars",
                   "config": "C5", "models": n_models, "subgrids": n_sub, "subgrids_per_gpu": L, "bands": nb,
                   "stars": stars, "threshold": THRESHOLD, "max_nbins": args.max_nbins,
                   "outputs": "merged stats+pdf1d", "parallelism": "subgrids sharded, all stars on every GPU, "
                   "per-batch all_gather + reduce_scatter by star",
                   "l2": "inputs larger than L2"},
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_stars * n_models / e2e_s, "unit": "star*model/s", "seconds_per_step": e2e_s,
                "stars": e2e_stars, "h2d_bytes_per_step": int(fluxes[:e2e_stars].nbytes),
                "d2h_bytes_per_step": int(rows * (st0.total_bins + 5 * st0.n_q + 3) * 8),
                "call": "SubgridFit.fit_batch per batch: host fluxes in, this rank's share of the merged rows "
                        "(pdf1d, percentiles, expectations, best values, Pmax, total_log_norm, chi2min) back on the host"},
        "gpu_launches": int(launches),
        "collective": {"payload_bytes_per_step_per_rank": int(coll_bytes), "ms_per_step": coll_ms,
                       "bus_gbs": (coll_bytes * (world - 1) / world / (coll_ms * 1e-3) / 1e9) if coll_ms and world > 1 else None,
                       "nvlink_gbs_reference": 770.0,
                       "note": "reduce_scatter(SUM) by star of weight-scaled pdf1d + packed expectation/best rows, "
                               "all_gather of total_log_norm/Pmax/chi2min; bus bandwidth = payload x (N-1)/N / device time"},
        "roofline": {"bound": "fp32", "kernel": "whole step", "achieved": (4 * nb + 2) * stars * n_models / world / (ms_step * 1e-3) / 1e12,
                     "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": (4 * nb + 2) * stars * n_models / world / (ms_step * 1e-3) / 1e12 / peak_tf, "traffic": None,
                     "frac_of_fp32_peak": (4 * nb + 2) * stars * n_models / world / (ms_step * 1e-3) / 1e12 / peak_tf},
        "stage_ms_per_step_rank0": stage_tot,
        "cpu_baseline": None, "setup_seconds": setup_s,
    }
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()


def main():
    # keep stdout for the one JSON line: libraries (NCCL's version banner, ...) write there too
    global _OUT
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.config == "C5":
        run_c5(args, rank, world, local_rank)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
