"""How much of the dense sweep a per-group lower bound could skip (CPU experiment, synthetic grids).

acc(s, m) = sum_b r_mb^2 (f_sb - m_mb)^2 + c2_m  >=  sum_b rlo_gb^2 dist(f_sb, [mlo_gb, mhi_gb])^2 + c2min_g
for every model m of group g.  A (star, group) pair can be skipped when that bound exceeds the
star's coarse minimum (every 32nd model) + tau + margin.  A quarter of a batch is simulated, so a
run of W/4 neighbouring stars here stands for W neighbouring stars of a full batch.
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from beast_b200 import synth  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
nstars = int(sys.argv[2]) if len(sys.argv) > 2 else 8448
scale = 33792 // nstars
t0 = time.time()
g, nz, obs = synth.make_config(cfg, nstars=nstars)
seds = np.asarray(g.seds)
m = torch.from_numpy((seds + nz["bias"]).astype(np.float64))
r = torch.from_numpy(1.0 / np.abs(nz["error"]))
w = np.asarray(g["weight"], dtype=np.float64)
nm, nb = m.shape
lnq = nb * 0.5 * np.log(2 * np.pi) - 0.5 * torch.log(r * r).sum(1).numpy()
c = -lnq + np.log(w)
c2 = torch.from_numpy(-2.0 * (c - c.max()))
print("grid", tuple(m.shape), "t=%.1f" % (time.time() - t0), flush=True)
f = torch.from_numpy(obs.fluxes.astype(np.float64))


def sweep(ms, rs, cs):
    out_min = torch.empty(nstars, dtype=torch.float64)
    out_arg = torch.empty(nstars, dtype=torch.int64)
    for a in range(0, nstars, 128):
        fb = f[a:a + 128]
        acc = cs[None, :].expand(len(fb), -1).clone()
        for b in range(nb):
            t = (fb[:, b, None] - ms[None, :, b]) * rs[None, :, b]
            acc.addcmul_(t, t)
        out_min[a:a + 128], out_arg[a:a + 128] = acc.min(dim=1)
    return out_min, out_arg


cache = "gpurun_out/prune_%s_%d.pt" % (cfg, nstars)
if os.path.exists(cache):
    cmin, carg, tmin = torch.load(cache)
else:
    cmin, carg = sweep(m[::32], r[::32], c2[::32])
    print("coarse done t=%.1f" % (time.time() - t0), flush=True)
    tmin = None
    if os.environ.get("TRUE_MIN"):
        tmin, _ = sweep(m, r, c2)
        print("full done t=%.1f" % (time.time() - t0), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    torch.save((cmin, carg, tmin), cache)

tau, margin = 20.0, 1.0


def bounds(G):
    ng = nm // G
    mg = m[:ng * G].reshape(ng, G, nb)
    rg = r[:ng * G].reshape(ng, G, nb)
    return mg.min(1).values, mg.max(1).values, rg.min(1).values, c2[:ng * G].reshape(ng, G).min(1).values


def lb(fs, bd):
    mlo, mhi, rlo, c2min = bd
    acc = c2min[None, :].expand(len(fs), -1).clone()
    for b in range(nb):
        u = torch.clamp(torch.maximum(mlo[None, :, b] - fs[:, b, None], fs[:, b, None] - mhi[None, :, b]), min=0.0)
        d = u * rlo[None, :, b]
        acc.addcmul_(d, d)
    return acc


def report(label, order, G, cut):
    bd = bounds(G)
    fo, co = f[order], cut[order]
    blk = 1024 // scale
    tot = {}
    n = 0
    for a in range(0, nstars - blk + 1, blk * max(1, nstars // blk // 6)):
        keep = lb(fo[a:a + blk], bd) <= (co[a:a + blk, None] + tau + margin)
        n += keep.numel()
        tot["star"] = tot.get("star", 0) + int(keep.sum())
        for wd in (32, 64, 256, 1024):
            k = keep.reshape(blk // (wd // scale), wd // scale, -1).any(dim=1)
            tot[wd] = tot.get(wd, 0) + int(k.sum()) * (wd // scale)
    print("%-40s G=%-3d" % (label, G), " ".join("%s=%.3f" % (k, v / n) for k, v in tot.items()),
          "t=%.0f" % (time.time() - t0), flush=True)


by_group = torch.argsort(carg // 2, stable=True)
by_model = torch.argsort(carg, stable=True)
lf = torch.log10(torch.clamp(f, min=1e-30))
by_flux0 = torch.argsort(lf.sum(1), stable=True)
report("coarse arg-min group (today)", by_group, 64, cmin)
report("coarse arg-min model", by_model, 64, cmin)
report("total log flux", by_flux0, 64, cmin)
report("coarse arg-min model", by_model, 32, cmin)
if tmin is not None:
    report("coarse arg-min model, true minimum", by_model, 64, tmin)

if os.environ.get("KEYS"):
    med = f.median(dim=0).values
    sl = torch.asinh(f / med)          # symmetric log of the flux in units of the band's median
    bright = sl.sum(1)
    report("asinh brightness", torch.argsort(bright, stable=True), 64, cmin)
    for b in range(nb):
        report("asinh flux band %d" % b, torch.argsort(sl[:, b], stable=True), 64, cmin)
    colour = sl[:, 0] - sl[:, nb - 1]

    def quant(x, n):
        r = torch.argsort(torch.argsort(x))
        return (r * n // len(x)).to(torch.int64)

    for nlev in (8, 32, 128):
        key = quant(bright, nlev) * (1 << 20) + quant(colour, 1 << 20)
        report("brightness in %d levels, then colour" % nlev, torch.argsort(key, stable=True), 64, cmin)
    for nlev in (8, 32):
        key = quant(colour, nlev) * (1 << 20) + quant(bright, 1 << 20)
        report("colour in %d levels, then brightness" % nlev, torch.argsort(key, stable=True), 64, cmin)
    key = (carg // 64) * (1 << 20) + quant(bright, 1 << 20)
    report("coarse group, then brightness", torch.argsort(key, stable=True), 64, cmin)

if os.environ.get("KEYS2"):
    med = f.median(dim=0).values
    sl = torch.asinh(f / med)

    def nested(primary, secondary, block):
        o = torch.argsort(sl[:, primary], stable=True)
        o = o.clone()
        for a in range(0, nstars, block):
            seg = o[a:a + block]
            o[a:a + block] = seg[torch.argsort(sl[seg, secondary], stable=True)]
        return o

    report("band 4", torch.argsort(sl[:, 4], stable=True), 64, cmin)
    report("band 3", torch.argsort(sl[:, 3], stable=True), 64, cmin)
    for blk in (64, 128, 256, 512):   # (a quarter of the real batch: 256 here = 1024 stars there)
        for sec in (2, 5, 3):
            report("band 4, blocks of %d by band %d" % (blk * scale, sec), nested(4, sec, blk), 64, cmin)
