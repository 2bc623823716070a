"""Subgrid mode across GPUs, checked end to end (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tools/subgrid_nccl_check.py [--axes small] [--stars 64]

Every rank fits its own row slice of the grid (do_not_normalize, shared grid_info_dict) on its own
GPU, the ranks merge with beast_b200.subgrid.merge_subgrids over NCCL, and rank 0 compares the
result with per-subgrid oracle fits + the restated merge arithmetic.  Prints one
JSON line with the largest deviations and the time of the NCCL merge."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--axes", default="small")
    ap.add_argument("--stars", type=int, default=64)
    ap.add_argument("--max-nbins", type=int, default=50)
    ap.add_argument("--no-oracle", action="store_true")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from beast_b200 import subgrid, synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)

    g = synth.make_sedgrid(args.axes, seed=31)
    nz = synth.make_noisemodel(g)
    obs = synth.make_observations(g, nz, args.stars, seed=32, with_coords=False)
    keys = [k for k in g.keys() if k not in
            "osl keep weight grid_weight prior_weight fullgrid_idx stage specgrid_indx phase eep".split()]
    slices = subgrid.uniform_slices(len(g), world)
    subs = [(g.subgrid(s), {k: v[s] for k, v in nz.items()}) for s in slices]
    info = subgrid.reduce_grid_info([subgrid.subgrid_info(sg, sn) for sg, sn in subs])
    sg, sn = subs[rank]
    setup = FitSetup(sg, sn, list(keys), max_nbins=args.max_nbins, grid_info_dict=info, do_not_normalize=True)
    eng = FitEngine(setup, device=dev)
    thr = -40.0
    res = eng.fit(obs.fluxes, threshold=thr)
    part = subgrid.local_result_tensors(setup, res, dev)
    math = subgrid.CudaMergeMath(setup)
    merged = subgrid.merge_subgrids(part, math, group=dist.group.WORLD, rank=rank, world=world)  # warm-up
    torch.cuda.synchronize(dev)
    dist.barrier()
    t0 = time.perf_counter()
    merged = subgrid.merge_subgrids(part, math, group=dist.group.WORLD, rank=rank, world=world)
    torch.cuda.synchronize(dev)
    merge_s = time.perf_counter() - t0
    out = {"ranks": world, "models": len(g), "stars": args.stars, "merge_seconds": merge_s}
    if rank == 0 and not args.no_oracle:
        from oracle import fit_oracle as fo

        oinfo = fo.reduce_grid_info([fo.subgrid_info(a, b) for a, b in subs])
        oparts = []
        for a, b in subs:
            ost = fo.OracleSetup(a, b, keys, max_nbins=args.max_nbins, grid_info_dict=oinfo, do_not_normalize=True)
            oparts.append(fo.fit_all(obs.fluxes, ost, thr, fast=True, keep_sparse=False))
        want = fo.merge_subgrid_results(oparts)
        m = {k: v.cpu().numpy() for k, v in merged.items()}
        out["owner_equal"] = bool(np.array_equal(m["best_gridsub_tag"], want["best_gridsub_tag"]))
        out["pmax_indx_equal"] = bool(np.array_equal(m["Pmax_indx"], want["Pmax_indx"]))
        out["best_equal"] = bool(np.array_equal(m["best_vals"], want["best_vals"]))
        out["max_abs_total_log_norm"] = float(np.max(np.abs(m["total_log_norm"] - want["total_log_norm"])))
        wpdf = np.concatenate([p[:-1] for p in want["pdf1d"]], axis=1)
        out["max_abs_pdf1d"] = float(np.max(np.abs(m["pdf1d"] - wpdf)))
        out["max_abs_percentile"] = float(np.max(np.abs(m["per_vals"] - want["per_vals"])))
        out["max_rel_chi2min"] = float(np.max(np.abs(m["chi2min"] - want["chi2min"]) / np.abs(want["chi2min"])))
        ok = (out["owner_equal"] and out["pmax_indx_equal"] and out["best_equal"]
              and out["max_abs_total_log_norm"] < 1e-4 and out["max_abs_pdf1d"] < 1e-9)
        out["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()
    if rank == 0 and not args.no_oracle and not out.get("ok", True):
        sys.exit(1)


if __name__ == "__main__":
    main()
