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


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES) + ["trim_tiny"]
    for n in names:
        if n.startswith("trim"):
            make_trim_case(n)
        else:
            make_case(n)
