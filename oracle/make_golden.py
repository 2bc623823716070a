"""
TEST INFRASTRUCTURE -- generate tests/golden/*.npz by running the REFERENCE itself.

Run in the build container (needs /root/reference):

    python -m oracle.make_golden

Each fixture holds the synthetic inputs (so it is self-contained and survives
changes to ``beast_b200.synth``), the call arguments, and every array the
unmodified reference ``Q_all_memory`` handed to its writers (captured through
``oracle/ref_shim.py``), plus the per-star sparse sets computed with the
reference's own ``N_logLikelihood_NM``.  The reference tree cannot travel to the
GPU box, these files can.
"""

from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from beast_b200 import synth  # noqa: E402
from oracle import fit_oracle, ref_shim  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

CASES = {
    # name: (axes, nstars, seed, kwargs, zero-weight stride)
    "tiny_thr10_pdf2d": ("tiny", 16, 11, dict(threshold=-10.0, max_nbins=50,
                                               pdf2d_param_list=["Av", "M_ini", "logT"]), 7),
    "tiny_thr40_nonorm": ("tiny", 12, 12, dict(threshold=-40.0, max_nbins=200,
                                               do_not_normalize=True), 0),
    "small_thr10": ("small", 24, 13, dict(threshold=-10.0, max_nbins=50), 5),
    # full-covariance (trunchen-like) noise model: N_covar_logLikelihood path, fit.py:432-444
    "tiny_fullcov": ("tiny", 12, 14, dict(threshold=-10.0, max_nbins=40), 6),
}


def pack_inputs(g, nz, obs):
    d = {"seds": g.seds, "lamb": g.lamb, "filters": np.array(g.filters),
         "colnames": np.array(g.keys()), "obs_flux": obs.fluxes}
    for k in g.keys():
        d["col__" + k] = g[k]
    for k in nz:
        d["noise__" + k] = nz[k]
    return d


def unpack_inputs(z):
    """rebuild (sedgrid, noisemodel, obs) from a fixture (used by the tests)"""
    filters = [str(f) for f in z["filters"]]
    cols = {str(k): z["col__" + str(k)] for k in z["colnames"]}
    g = synth.SyntheticSEDGrid(z["lamb"], z["seds"], cols, filters)
    nz = {k[len("noise__"):]: z[k] for k in z.files if k.startswith("noise__")}
    obs = synth.SyntheticObservations(z["obs_flux"], filters)
    return g, nz, obs


def make_case(name):
    axes, nstars, seed, kw, zstride = CASES[name]
    g = synth.make_sedgrid(axes, seed=seed)
    if zstride:
        g["weight"][::zstride] = 0.0
    nz = synth.make_noisemodel_fullcov(g) if "fullcov" in name else synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, nstars, seed=seed + 100)
    keys = fit_oracle.default_keys(g)
    ref = ref_shim.run_reference_fit(obs, g, nz, keys, **kw)
    out = pack_inputs(g, nz, obs)
    out["keys"] = np.array(keys)
    out["threshold"] = kw["threshold"]
    out["max_nbins"] = kw["max_nbins"]
    out["do_not_normalize"] = bool(kw.get("do_not_normalize", False))
    out["pdf2d_param_list"] = np.array(kw.get("pdf2d_param_list") or [], dtype=str)
    s = ref["stats"]
    out["ref_qnames"] = np.array(s["qnames"])
    for k in ("best_vals", "exp_vals", "per_vals", "chi2min", "chi2min_indx", "Pmax",
              "Pmax_indx", "specgrid_indx", "total_log_norm"):
        out["ref_" + k] = s[k]
    for k, a in enumerate(ref["pdf1d"]):
        out["ref_pdf1d_%03d" % k] = a
    out["ref_pdf2d_pairs"] = np.array(ref.get("pdf2d_pairs", []), dtype=str)
    for k, a in enumerate(ref.get("pdf2d", [])):
        out["ref_pdf2d_%03d" % k] = a
    off = [0]
    idx_all, lnp_all, chi2_all = [], [], []
    for e in range(nstars):
        idx, lnp, chi2 = ref_shim.reference_sparse_set(
            obs.fluxes[e], g, nz, kw["threshold"], bool(kw.get("do_not_normalize", False)))
        idx_all.append(idx)
        lnp_all.append(lnp)
        chi2_all.append(chi2)
        off.append(off[-1] + len(idx))
    out["ref_sparse_off"] = np.array(off, dtype=np.int64)
    out["ref_sparse_idx"] = np.concatenate(idx_all)
    out["ref_sparse_lnp"] = np.concatenate(lnp_all)
    out["ref_sparse_chi2"] = np.concatenate(chi2_all)
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.0f kB" % (os.path.getsize(path) / 1e3))


def make_trim_case(name="trim_tiny"):
    """inputs + the index list the reference's trim_models keeps (trim_grid.py:53-125)"""
    g = synth.make_sedgrid("tiny", seed=21)
    nz = synth.make_noisemodel(g)
    nz["completeness"][::13, 1] = 0.0
    obs = synth.make_observations(g, nz, 30, seed=22)
    bright = np.argsort(obs.fluxes[:, 3])[12:]
    obs = synth.SyntheticObservations(obs.fluxes[bright], g.filters)
    cap = ref_shim.run_reference_trim(g, nz, obs)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    np.savez_compressed(path, seds=g.seds, bias=nz["bias"], error=nz["error"],
                        completeness=nz["completeness"], obs_flux=obs.fluxes,
                        filters=np.array(g.filters), ref_indxs=cap["cols"]["fullgrid_idx"],
                        ref_seds=cap["seds"], ref_error=cap["noise_error"])
    print("wrote", path, "%.0f kB" % (os.path.getsize(path) / 1e3), len(cap["cols"]["fullgrid_idx"]), "of", len(g))


def stats_table(s, qnames, p=(16, 50, 84)):
    """the stats columns as the reference's save_stats names them (fit.py:94-108)"""
    t = {}
    for k, q in enumerate(qnames):
        t[q + "_Best"] = s["best_vals"][:, k]
        t[q + "_Exp"] = s["exp_vals"][:, k]
        for i, pv in enumerate(p):
            t["%s_p%d" % (q, int(pv))] = s["per_vals"][:, k, i]
    t["chi2min"] = s["chi2min"]
    t["chi2min_indx"] = s["chi2min_indx"].astype(int)
    t["Pmax"] = s["Pmax"]
    t["Pmax_indx"] = s["Pmax_indx"].astype(int)
    t["specgrid_indx"] = s["specgrid_indx"].astype(int)
    t["total_log_norm"] = s["total_log_norm"]
    return t


def make_merge_case(name="merge_tiny", n_sub=3, nstars=10, max_nbins=40, threshold=-10.0):
    """Subgrid mode end to end with the reference's own code (docs/subgrid_parallelism.rst:77-143):
    row slices of a grid, ``reduce_grid_info`` over them, one ``Q_all_memory`` per slice with
    ``do_not_normalize=True`` and the shared ``grid_info_dict``, then ``merge_pdf1d_stats`` and
    ``merge_lnp`` (subgridding_tools.py:189-271, 274-539, 542-658), all unmodified under the
    in-memory stand-ins of ``oracle/ref_shim.py``."""
    from beast_b200.subgrid import uniform_slices

    g = synth.make_sedgrid("tiny", seed=31)
    g["weight"][::9] = 0.0
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, nstars, seed=32)
    keys = fit_oracle.default_keys(g)
    slices = uniform_slices(len(g), n_sub)
    subs = [g.subgrid(s) for s in slices]
    nzs = [{k: v[s] for k, v in nz.items()} for s in slices]
    info = ref_shim.run_reference_grid_info(subs, nzs)
    out = pack_inputs(g, nz, obs)
    out["keys"] = np.array(keys)
    out["n_sub"], out["max_nbins"], out["threshold"] = n_sub, max_nbins, threshold
    out["info_keys"] = np.array(list(info.keys()))
    out["info_min"] = np.array([info[q]["min"] for q in info], dtype=float)
    out["info_max"] = np.array([info[q]["max"] for q in info], dtype=float)
    out["info_num_unique"] = np.array([info[q]["num_unique"] for q in info])
    for q in info:
        out["info_unique__" + q] = np.asarray(info[q]["unique_vals"])
    pdfs, tabs, lnps = [], [], []
    for gi, (sg, snz) in enumerate(zip(subs, nzs)):
        ref = ref_shim.run_reference_fit(obs, sg, snz, keys, threshold=threshold, max_nbins=max_nbins,
                                         grid_info_dict=info, do_not_normalize=True)
        qnames = ref["stats"]["qnames"]
        pdfs.append(ref["pdf1d"])
        tabs.append(stats_table(ref["stats"], qnames))
        stars = []
        for e in range(nstars):
            idx, lnp, _ = ref_shim.reference_sparse_set(obs.fluxes[e], sg, snz, threshold, True)
            stars.append((idx, lnp.astype(np.float32).astype(float)))  # stored float32 (fit.py:705)
        lnps.append(stars)
        for k, a in enumerate(ref["pdf1d"]):
            out["sub%d_pdf1d_%03d" % (gi, k)] = a
        for c, v in tabs[-1].items():
            out["sub%d_stats__%s" % (gi, c)] = np.asarray(v)
    out["ref_qnames"] = np.array(qnames)
    m_pdf, m_stats = ref_shim.run_reference_merge_pdf1d_stats(pdfs, qnames, tabs)
    for k, a in enumerate(m_pdf):
        out["merged_pdf1d_%03d" % k] = a
    for c, v in m_stats.items():
        out["merged_stats__" + c] = np.asarray(v)
    m_lnp = ref_shim.run_reference_merge_lnp(lnps, threshold=threshold)
    for e, d in enumerate(m_lnp):
        for c in ("idx", "lnp", "subgrid"):
            out["merged_lnp_%03d_%s" % (e, c)] = np.asarray(d[c], dtype=float)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.0f kB" % (os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES) + ["trim_tiny", "merge_tiny"]
    for n in names:
        if n.startswith("trim"):
            make_trim_case(n)
        elif n.startswith("merge"):
            make_merge_case(n)
        else:
            make_case(n)
