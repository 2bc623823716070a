"""The N>1 choreography of the subgrid merge on CPU: two gloo ranks, each holding the
(oracle-computed) results of one subgrid, run beast_b200.subgrid.merge_subgrids with its
collectives; the arithmetic hooks are a NumPy stand-in defined here (the product's own are
CUDA kernels).  Result must equal the single-process restated merge."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import fit_oracle as fo


class NumpyMergeMath:
    """test double for beast_b200.subgrid.CudaMergeMath (same three hooks)"""

    def __init__(self, bins):
        self.bins = bins

    def weights(self, tln_all):
        t = tln_all.numpy()
        mx = t.max(axis=0)
        w = np.exp(t - mx[None, :])
        return torch.from_numpy(w), torch.from_numpy(np.log(w.sum(axis=0)) + mx)

    def scale_rows(self, rows, w):
        return rows * w.reshape(-1, *([1] * (rows.dim() - 1)))

    def finalize(self, pdf):
        p = pdf.numpy()
        per = np.zeros((p.shape[0], len(self.bins), 3))
        off = 0
        for k, b in enumerate(self.bins):
            blk = p[:, off:off + b.nbins]
            norm = blk.sum(axis=1, keepdims=True)
            np.divide(blk, norm, out=blk, where=norm > 0)
            for e in range(p.shape[0]):
                if blk[e].max() > 0:
                    per[e, k] = fo.percentile(b.bin_vals, [16.0, 50.0, 84.0], blk[e])
            off += b.nbins
        return torch.from_numpy(per)


def _make_parts():
    from beast_b200 import subgrid, synth

    g = synth.make_sedgrid("tiny", seed=41)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 9, seed=42)
    keys = fo.default_keys(g)
    slices = subgrid.uniform_slices(len(g), 2)
    subs = [(g.subgrid(s), {k: v[s] for k, v in nz.items()}) for s in slices]
    info = fo.reduce_grid_info([fo.subgrid_info(sg, sn) for sg, sn in subs])
    outs, setups = [], []
    for sg, sn in subs:
        ost = fo.OracleSetup(sg, sn, keys, max_nbins=30, grid_info_dict=info, do_not_normalize=True)
        outs.append(fo.fit_all(obs.fluxes, ost, -40.0, fast=True, keep_sparse=False))
        setups.append(ost)
    return outs, setups


def _as_local(o):
    return {
        "total_log_norm": torch.from_numpy(o["total_log_norm"]), "Pmax": torch.from_numpy(o["Pmax"]),
        "Pmax_indx": torch.from_numpy(o["Pmax_indx"].astype(np.int64)),
        "chi2min": torch.from_numpy(o["chi2min"]), "best_vals": torch.from_numpy(o["best_vals"]),
        "exp_vals": torch.from_numpy(o["exp_vals"]),
        "pdf1d": torch.from_numpy(np.concatenate([p[:-1] for p in o["pdf1d"]], axis=1)),
    }


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from beast_b200 import subgrid

        outs, setups = _make_parts()
        merged = subgrid.merge_subgrids(_as_local(outs[rank]), NumpyMergeMath(setups[0].bins),
                                        group=dist.group.WORLD, rank=rank, world=world)
        q.put((rank, {k: v.numpy() for k, v in merged.items()}))
    finally:
        dist.destroy_process_group()


def test_two_rank_merge_equals_single_process_merge():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    outs, setups = _make_parts()
    want = fo.merge_subgrid_results(outs)
    for rank in (0, 1):
        m = got[rank]
        np.testing.assert_array_equal(m["best_gridsub_tag"], want["best_gridsub_tag"])
        np.testing.assert_array_equal(m["Pmax_indx"], want["Pmax_indx"])
        np.testing.assert_array_equal(m["best_vals"], want["best_vals"])
        np.testing.assert_allclose(m["Pmax"], want["Pmax"], rtol=0, atol=0)
        np.testing.assert_allclose(m["chi2min"], want["chi2min"], rtol=0, atol=0)
        np.testing.assert_allclose(m["total_log_norm"], want["total_log_norm"], rtol=1e-14)
        np.testing.assert_allclose(m["exp_vals"], want["exp_vals"], rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(m["pdf1d"], np.concatenate([p[:-1] for p in want["pdf1d"]], axis=1),
                                   rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(m["per_vals"], want["per_vals"], rtol=1e-10, atol=1e-300)


def test_single_process_list_merge_uses_same_path():
    from beast_b200 import subgrid

    outs, setups = _make_parts()
    merged = subgrid.merge_subgrids([_as_local(o) for o in outs], NumpyMergeMath(setups[0].bins))
    want = fo.merge_subgrid_results(outs)
    np.testing.assert_array_equal(merged["best_gridsub_tag"].numpy(), want["best_gridsub_tag"])
    np.testing.assert_allclose(merged["total_log_norm"].numpy(), want["total_log_norm"], rtol=1e-14)


# ---- per-batch device merge (SubgridFit): 2 ranks x 2 resident subgrids each -----------------
class _Setup:
    """what SubgridFit needs of a FitSetup: identity map to subgrid rows + parameter columns"""

    def __init__(self, ost):
        self.g0_indxs = np.arange(ost.nmodels, dtype=np.int64)
        self.qcols = [np.asarray(q, dtype=np.float64) for q in ost.qvals]


def _four_parts():
    from beast_b200 import subgrid, synth

    g = synth.make_sedgrid("tiny", seed=51)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, 7, seed=52)  # (7 stars: the shares of 2 ranks are uneven)
    keys = fo.default_keys(g)
    subs = [(g.subgrid(s), {k: v[s] for k, v in nz.items()}) for s in subgrid.uniform_slices(len(g), 4)]
    info = fo.reduce_grid_info([fo.subgrid_info(sg, sn) for sg, sn in subs])
    outs, setups = [], []
    for sg, sn in subs:
        ost = fo.OracleSetup(sg, sn, keys, max_nbins=30, grid_info_dict=info, do_not_normalize=True)
        outs.append(fo.fit_all(obs.fluxes, ost, -40.0, fast=True, keep_sparse=False))
        setups.append(ost)
    return outs, setups


def _as_batch(o):
    """an oracle subgrid result dressed as the device dict FitEngine.fit_batch returns"""
    return {"total_log_norm": torch.from_numpy(o["total_log_norm"]), "Pmax": torch.from_numpy(o["Pmax"]),
            "chi2min": torch.from_numpy(o["chi2min"]),
            "Pmax_j": torch.from_numpy(o["Pmax_indx"].astype(np.int64)),
            "exp_vals": torch.from_numpy(o["exp_vals"]),
            "pdf1d": torch.from_numpy(np.concatenate([p[:-1] for p in o["pdf1d"]], axis=1))}


def _fit_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from beast_b200 import subgrid
        from tests.util import NumpyMergeMath as Math

        outs, setups = _four_parts()
        mine = [2 * rank, 2 * rank + 1]
        sf = subgrid.SubgridFit([None, None], Math([b.bin_vals for b in setups[0].bins]),
                                group=dist.group.WORLD, rank=rank, world=world)
        m = sf.merge_batch([_as_batch(outs[i]) for i in mine], [_Setup(setups[i]) for i in mine])
        q.put((rank, {k: (v.numpy() if hasattr(v, "numpy") else v) for k, v in m.items()}))
    finally:
        dist.destroy_process_group()


def test_per_batch_merge_over_two_ranks_with_two_subgrids_each():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_fit_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    outs, setups = _four_parts()
    want = fo.merge_subgrid_results(outs)
    assert got[0]["rows"] == (0, 4) and got[1]["rows"] == (4, 7)
    for rank in (0, 1):
        m = got[rank]
        lo, hi = m["rows"]
        np.testing.assert_array_equal(m["best_gridsub_tag"], want["best_gridsub_tag"][lo:hi])
        np.testing.assert_array_equal(m["Pmax_indx"], want["Pmax_indx"][lo:hi])
        np.testing.assert_array_equal(m["best_vals"], want["best_vals"][lo:hi])
        np.testing.assert_array_equal(m["Pmax"], want["Pmax"][lo:hi])
        np.testing.assert_array_equal(m["chi2min"], want["chi2min"][lo:hi])
        np.testing.assert_allclose(m["total_log_norm"], want["total_log_norm"][lo:hi], rtol=1e-14)
        np.testing.assert_allclose(m["exp_vals"], want["exp_vals"][lo:hi], rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(m["pdf1d"], np.concatenate([p[lo:hi] for p in want["pdf1d"]], axis=1),
                                   rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(m["per_vals"], want["per_vals"][lo:hi], rtol=1e-10, atol=1e-300)
