"""Where the host time of one Q_all_memory call goes (C2, engine with a resident grid): cProfile
of the third call, plus the wall time of the call against the device time of its kernels."""
import contextlib
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from beast_b200 import fit as bfit  # noqa: E402
from beast_b200.engine import FitEngine  # noqa: E402
from beast_b200.gridprep import FitSetup  # noqa: E402

stars = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
g, nz, obs = bench.build_inputs("C2", stars, seed=4321)
keys = [k for k in g.keys() if k not in
        "osl keep weight grid_weight prior_weight fullgrid_idx stage specgrid_indx phase eep".split()]
setup = FitSetup(g, nz, keys, max_nbins=200)
eng = FitEngine(setup, device=torch.device("cuda", 0), gmin_budget_bytes=24 << 30)
eng.reuse_buffers = True
kw = dict(p=[16.0, 50.0, 84.0], max_nbins=200, lnp_npts=500, threshold=-10.0, engine=eng, keep_lnp=True)


def call():
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        return bfit.Q_all_memory({}, obs, g, nz, list(keys), **kw)


call()
call()
torch.cuda.synchronize()
t0 = time.perf_counter()
call()
print("wall of one call: %.1f ms" % ((time.perf_counter() - t0) * 1e3))
pr = cProfile.Profile()
pr.enable()
call()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(28)
