"""
Seeded synthetic inputs for the fitting hot path.

Isochrone / atmosphere / filter libraries cannot be downloaded offline, so the
SED grid, the noise model and the star catalog are generated analytically.  The
grid is a *smooth tensor product* (logA x M_ini x Z x Av x Rv x f_A [x distance])
so that neighbouring models are nearly degenerate and the per-star sparse sets
have realistic sizes -- a random grid would give one-model posteriors and leave
the histogram / percentile stage idle.

The objects returned here duck-type exactly the surface the reference fit
touches (reference ``beast/fitting/fit.py:396-463, 624-637, 858-913``):

* ``SyntheticSEDGrid``  -- ``.seds``, ``.lamb``, ``.filters``, ``[name]``, ``.keys()``
  (column names follow ``docs/beast_grid_format.rst:24-78``)
* noise model           -- plain ``dict`` with ``error``, ``bias``, ``completeness``
  (``observationmodel/noisemodel/generic_noisemodel.py:110-139``)
* ``SyntheticObservations`` -- ``len()``, ``enumobs()``, ``getFilters()``,
  ``.data``, ``.filters``, ``.filter_aliases``, ``.vega_flux``
  (``observationmodel/observations.py:120-210``)

Stars are drawn the way ``gen_SimObs_from_sedgrid`` draws them
(``observationmodel/observations.py:402-422``): model index with probability
proportional to ``weight x completeness``, flux = seds + bias + N(0, error).
"""

from __future__ import annotations

import numpy as np

__all__ = [
    "PHAT_FILTERS",
    "UVNIR_FILTERS",
    "CONFIG_AXES",
    "SyntheticSEDGrid",
    "SyntheticObservations",
    "make_sedgrid",
    "make_noisemodel",
    "make_noisemodel_fullcov",
    "make_observations",
    "make_config",
]

# filter names as they appear in docs/outputs.rst:35-40 of the reference
PHAT_FILTERS = [
    "HST_WFC3_F275W",
    "HST_WFC3_F336W",
    "HST_ACS_WFC_F475W",
    "HST_ACS_WFC_F814W",
    "HST_WFC3_F110W",
    "HST_WFC3_F160W",
]
PHAT_LAMB = np.array([2750.0, 3360.0, 4750.0, 8140.0, 11000.0, 16000.0])

# 8-band UV-NIR set (plotting/tests/test_plots.py:59-68 of the reference)
UVNIR_FILTERS = [
    "HST_WFC3_F225W",
    "HST_WFC3_F275W",
    "HST_WFC3_F336W",
    "HST_ACS_WFC_F475W",
    "HST_ACS_WFC_F550M",
    "HST_ACS_WFC_F814W",
    "HST_WFC3_F110W",
    "HST_WFC3_F160W",
]
UVNIR_LAMB = np.array([2250.0, 2750.0, 3360.0, 4750.0, 5580.0, 8140.0, 11000.0, 16000.0])

# (n_logA, n_M, n_Z, n_Av, n_Rv, n_fA, n_dist) per named configuration
CONFIG_AXES = {
    "tiny": (4, 8, 2, 5, 3, 2, 1),  # 1,920 models (unit tests / golden vectors)
    "small": (6, 14, 3, 8, 4, 2, 1),  # 16,128 models
    "C1": (10, 25, 4, 10, 5, 2, 1),  # 1e5 models
    "C2": (50, 80, 5, 20, 5, 2, 1),  # 4e6 models
    "C4": (50, 100, 5, 20, 5, 4, 1),  # 1e7 models
    "C5": (50, 100, 5, 20, 5, 4, 2),  # 2e7 models (split as 8 subgrids)
}

_H = 6.62607015e-27
_C = 2.99792458e10
_KB = 1.380649e-16
_RSUN = 6.957e10
_PC = 3.0856775814913673e18


def _planck_lambda(lamb_A, T):
    """pi*B_lambda in erg/s/cm^2/A for wavelength in Angstrom, T in K."""
    lam = lamb_A * 1e-8  # cm
    x = _H * _C / (lam * _KB * T)
    b = 2.0 * _H * _C**2 / lam**5 / np.expm1(x)  # per cm
    return np.pi * b * 1e-8  # per Angstrom


def _ext_curve(lamb_A, Rv, f_A):
    """smooth A_lambda/A_V mixture curve (analytic stand-in for the dust model)"""
    x = 5500.0 / lamb_A
    k_a = x ** (0.7 + 2.0 / Rv)
    k_b = x**1.3 * (1.0 + 0.25 * (x - 1.0))
    return f_A * k_a + (1.0 - f_A) * k_b


class SyntheticSEDGrid:
    """In-memory stand-in for ``beast.physicsmodel.grid.SEDGrid``.

    Only the members the fit path touches are provided.
    """

    def __init__(self, lamb, seds, cols, filters):
        self.lamb = np.asarray(lamb, dtype=float)
        self.seds = np.asarray(seds, dtype=float)
        self.filters = list(filters)
        self._cols = dict(cols)
        self.header = {"filters": " ".join(self.filters)}

    def keys(self):
        return list(self._cols.keys())

    def __getitem__(self, name):
        return self._cols[name]

    def __contains__(self, name):
        return name in self._cols

    def __len__(self):
        return self.seds.shape[0]

    def subgrid(self, slc):
        """row slice, like ``split_grid`` does (tools/subgridding_tools.py:66-84)"""
        return SyntheticSEDGrid(
            self.lamb,
            self.seds[slc],
            {k: v[slc] for k, v in self._cols.items()},
            self.filters,
        )


class SyntheticObservations:
    """Duck-typed ``Observations``: fluxes are stored directly in cgs units."""

    def __init__(self, fluxes, filters, ra=None, dec=None, extra=None):
        self.fluxes = np.ascontiguousarray(fluxes, dtype=float)
        self.filters = list(filters)
        self.nObs = self.fluxes.shape[0]
        self.vega_flux = np.ones(len(self.filters))
        self.filter_aliases = {f: f.split("_")[-1] + "_RATE" for f in self.filters}
        self.data = {}
        for k, f in enumerate(self.filters):
            self.data[self.filter_aliases[f]] = self.fluxes[:, k]
        if ra is not None:
            self.data["RA"] = np.asarray(ra, dtype=float)
            self.data["DEC"] = np.asarray(dec, dtype=float)
        if extra:
            self.data.update(extra)

    def __len__(self):
        return self.nObs

    def getFilters(self):
        return self.filters

    def getObs(self, num=0):
        return self.fluxes[num].copy()

    def enumobs(self):
        for k in range(self.nObs):
            yield k, self.getObs(k)

    def subset(self, slc):
        extra = {
            k: v[slc]
            for k, v in self.data.items()
            if k not in self.filter_aliases.values() and k not in ("RA", "DEC")
        }
        ra = self.data["RA"][slc] if "RA" in self.data else None
        dec = self.data["DEC"][slc] if "DEC" in self.data else None
        return SyntheticObservations(self.fluxes[slc], self.filters, ra, dec, extra)


def make_sedgrid(axes="C1", filters=None, lamb=None, seed=1234, rows=None):
    """Build the tensor-product SED grid.

    ``axes`` is a name from ``CONFIG_AXES`` or an explicit 7-tuple.  Row order is
    C-order over (logA, M_ini, Z, Av, Rv, f_A, distance) -- stellar parameters
    outermost, dust innermost, which is how BEAST grids are laid out.

    ``rows=(r0, r1)``: only that row slice of the grid (a subgrid as ``split_grid`` cuts them,
    tools/subgridding_tools.py:39-91), built without materialising the rest; weights are
    normalised over the whole grid, and the seeded roughness is drawn per slice, so a slice is
    not bit-identical to the same rows of the full grid.
    """
    if isinstance(axes, str):
        axes = CONFIG_AXES[axes]
    n_logA, n_M, n_Z, n_Av, n_Rv, n_fA, n_d = axes
    if filters is None:
        filters, lamb = PHAT_FILTERS, PHAT_LAMB
    lamb = np.asarray(lamb, dtype=float)
    nb = len(filters)

    a_logA = np.linspace(6.0, 10.13, n_logA)
    a_M = np.logspace(np.log10(0.35), np.log10(80.0), n_M)
    a_Z = np.array([0.0004, 0.004, 0.008, 0.019, 0.03, 0.05])[:n_Z] if n_Z <= 6 else np.logspace(-3.4, -1.3, n_Z)
    a_Av = np.linspace(0.0, 10.0, n_Av) if n_Av > 1 else np.array([0.0])
    a_Rv = np.linspace(2.0, 6.0, n_Rv) if n_Rv > 1 else np.array([3.1])
    a_fA = np.linspace(0.0, 1.0, n_fA) if n_fA > 1 else np.array([1.0])
    a_d = 783e3 * np.linspace(1.0, 1.25, n_d) if n_d > 1 else np.array([783e3])

    # ---- stellar part: one entry per (logA, M, Z) = the "spectral grid" ----
    gA, gM, gZ = np.meshgrid(a_logA, a_M, a_Z, indexing="ij")
    lgM = np.log10(gM)
    t_ms = 10.0 - 2.5 * lgM  # log10 main-sequence lifetime [yr]
    x = 10.0 ** np.clip(gA - t_ms, -6.0, 0.6)  # age / t_ms
    evol = 1.0 / (1.0 + np.exp(-8.0 * (x - 1.0)))  # smooth MS -> giant switch
    logL = 3.6 * lgM - 0.35 * lgM**2 + 0.45 * np.minimum(x, 1.5) + 0.9 * evol
    logT = 3.762 + 0.58 * lgM - 0.06 * lgM**2 - 0.02 * np.log10(gZ / 0.019) - 0.03 * np.minimum(x, 1.5) - 0.22 * evol
    logR = 0.5 * logL - 2.0 * (logT - 3.762)  # log R/Rsun
    logg = lgM - 2.0 * logR + 4.438
    M_act = gM * (1.0 - 0.12 * np.minimum(x, 1.0) - 0.2 * evol)
    mbol = 4.77 - 2.5 * logL

    # mass prior (Kroupa-like) x grid spacing, flat in the rest
    dM = np.gradient(a_M)
    dA = np.gradient(10.0**a_logA) if n_logA > 1 else np.ones(1)
    gw_star = (dA[:, None, None] / dA.sum()) * (dM[None, :, None] / dM.sum()) * np.ones_like(gZ)
    pw_star = gM**-2.3
    pw_star = pw_star / pw_star.mean()

    T = 10.0**logT
    R = 10.0**logR * _RSUN
    n_star_all = gA.size
    n_dust_all = n_Av * n_Rv * n_fA * n_d
    w_total = float(((gw_star / n_dust_all) * pw_star).sum() * n_dust_all)
    nm_all = n_star_all * n_dust_all
    s0, s1 = 0, n_star_all
    if rows is not None:
        r0, r1 = int(rows[0]), int(rows[1])
        s0, s1 = r0 // n_dust_all, -(-r1 // n_dust_all)

    def _cut(a):
        return a.reshape(-1)[s0:s1]

    (gA, gM, gZ, logL, logT, logR, logg, M_act, mbol, gw_star, pw_star, T, R) = [
        _cut(a) for a in (gA, gM, gZ, logL, logT, logR, logg, M_act, mbol, gw_star, pw_star, T, R)]
    n_star = gA.size
    f0 = _planck_lambda(lamb[None, :], T.reshape(-1, 1)) * (R.reshape(-1, 1) ** 2)  # x 1/d^2 later

    # ---- dust part ----
    gAv, gRv, gfA, gD = np.meshgrid(a_Av, a_Rv, a_fA, a_d, indexing="ij")
    n_dust = gAv.size
    k = _ext_curve(lamb[None, :], gRv.reshape(-1, 1), gfA.reshape(-1, 1))
    att = 10.0 ** (-0.4 * gAv.reshape(-1, 1) * k)  # (n_dust, nb)
    inv_d2 = 1.0 / (gD.reshape(-1) * _PC) ** 2
    inv_d2_ref = 1.0 / (a_d[0] * _PC) ** 2

    nm = n_star * n_dust
    seds = (f0[:, None, :] * (att * inv_d2[:, None])[None, :, :]).reshape(nm, nb)
    seds_nd = np.broadcast_to((f0 * inv_d2_ref)[:, None, :], (n_star, n_dust, nb)).reshape(nm, nb)

    def star_col(v):
        return np.repeat(v.reshape(-1), n_dust)

    def dust_col(v):
        return np.tile(v.reshape(-1), n_star)

    cols = {}
    cols["Av"] = dust_col(gAv)
    cols["Rv"] = dust_col(gRv)
    cols["f_A"] = dust_col(gfA)
    cols["Rv_A"] = 1.0 / (cols["f_A"] / cols["Rv"] + (1.0 - cols["f_A"]) / 2.74)
    cols["M_ini"] = star_col(gM)
    cols["logA"] = star_col(gA)
    cols["Z"] = star_col(gZ)
    cols["distance"] = dust_col(gD)
    cols["logL"] = star_col(logL)
    cols["logT"] = star_col(logT)
    cols["logg"] = star_col(logg)
    cols["M_act"] = star_col(M_act)
    cols["mbolmag"] = star_col(mbol)
    cols["radius"] = star_col(10.0**logR)
    for b, f in enumerate(filters):
        cols["log" + f + "_nd"] = np.log10(seds_nd[:, b])
        cols["log" + f + "_wd"] = np.log10(seds[:, b])
    cols["grid_weight"] = star_col(gw_star) / n_dust
    cols["prior_weight"] = star_col(pw_star)
    w = cols["grid_weight"] * cols["prior_weight"]
    cols["weight"] = w / w_total * nm_all
    cols["specgrid_indx"] = star_col(np.arange(s0, s1, dtype=np.int64))

    # a hair of seeded roughness so ties between models are not systematic
    rng = np.random.default_rng(seed if rows is None else [seed, s0])
    seds = seds * (1.0 + 1e-3 * rng.standard_normal(seds.shape))
    for b, f in enumerate(filters):
        cols["log" + f + "_wd"] = np.log10(seds[:, b])

    grid = SyntheticSEDGrid(lamb, seds, cols, filters)
    if rows is not None:
        grid = grid.subgrid(slice(r0 - s0 * n_dust_all, r1 - s0 * n_dust_all))
    return grid


def make_noisemodel(sedgrid, snr_cap=100.0, depth_scale=1.0):
    """Toothpick-like noise model: independent bands, per-model bias and sigma.

    error = sqrt((flux/snr_cap)^2 + floor_b^2); bias saturates at 1% of the floor
    (crowding); completeness is a smooth function of the best-band S/N, floored
    above zero so that the fit's completeness check passes (fit.py:404-409).
    """
    seds = sedgrid.seds
    nb = seds.shape[1]
    lamb = sedgrid.lamb
    # per-band noise floor: deeper in the optical, shallower in the UV/NIR
    floor = depth_scale * 2.0e-20 * (1.0 + 3.0 * np.abs(np.log10(lamb / 6000.0))) ** 1.5
    floor = floor[:nb]
    error = np.sqrt((seds / snr_cap) ** 2 + floor[None, :] ** 2)
    bias = 0.05 * floor[None, :] * np.tanh(seds / floor[None, :])
    snr = (seds / error).max(axis=1)
    compl1 = 1.0 / (1.0 + np.exp(-(snr - 5.0) / 0.8))
    completeness = np.repeat(np.clip(compl1, 1e-4, 1.0)[:, None], nb, axis=1)
    return {"error": error, "bias": bias, "completeness": completeness}


def make_noisemodel_fullcov(sedgrid, snr_cap=100.0, depth_scale=1.0):
    """Trunchen-like noise model: the toothpick sigma/bias/completeness plus a full covariance
    per model, C = diag(sigma) R diag(sigma) with a band-to-band correlation that grows for
    faint (crowding-limited) models.  Adds the three arrays that switch the fit to the
    full-covariance likelihood (fit.py:432-444): ``q_norm`` = -1/2 ln det C,
    ``icov_diag`` and ``icov_offdiag`` (upper triangle of C^-1 packed row-major, k < l), the
    packing of ``observationmodel/noisemodel/trunchen.py:388-421``."""
    nz = make_noisemodel(sedgrid, snr_cap=snr_cap, depth_scale=depth_scale)
    sig = nz["error"]
    nm, nb = sig.shape
    snr = (sedgrid.seds / sig).max(axis=1)
    rho = 0.15 + 0.35 / (1.0 + np.exp((snr - 8.0) / 3.0))  # (nm,)
    lag = np.abs(np.arange(nb)[:, None] - np.arange(nb)[None, :])
    R = rho[:, None, None] ** lag[None, :, :]
    C = sig[:, :, None] * R * sig[:, None, :]
    iC = np.linalg.inv(C)
    _, logdet = np.linalg.slogdet(C)
    iu = np.triu_indices(nb, k=1)
    nz["q_norm"] = -0.5 * logdet
    nz["icov_diag"] = np.ascontiguousarray(np.diagonal(iC, axis1=1, axis2=2))
    nz["icov_offdiag"] = np.ascontiguousarray(iC[:, iu[0], iu[1]])
    return nz


def make_observations(sedgrid, noisemodel, nstars, seed=4321, with_coords=True):
    """Draw ``nstars`` simulated stars (observations.py:402-422 of the reference)."""
    rng = np.random.default_rng(seed)
    nm = len(sedgrid)
    p = sedgrid["weight"] * noisemodel["completeness"].max(axis=1)
    p = p / p.sum()
    sim_indx = rng.choice(nm, size=nstars, p=p)
    loc = sedgrid.seds[sim_indx] + noisemodel["bias"][sim_indx]
    flux = rng.normal(loc=loc, scale=np.abs(noisemodel["error"][sim_indx]))
    ra = dec = None
    if with_coords:
        ra = 10.68 + 0.5 * rng.random(nstars)
        dec = 41.27 + 0.5 * rng.random(nstars)
    obs = SyntheticObservations(flux, sedgrid.filters, ra, dec)
    obs.sim_indx = sim_indx
    return obs


def make_config(name, nstars=None, seed=1234):
    """(sedgrid, noisemodel, obs) for one of the BASELINE.json configurations."""
    default_stars = {"tiny": 16, "small": 64, "C1": 1000, "C2": 100000, "C4": 200000, "C5": 1000000}
    if name == "C4":
        g = make_sedgrid(name, UVNIR_FILTERS, UVNIR_LAMB, seed=seed)
    else:
        g = make_sedgrid(name, seed=seed)
    nz = make_noisemodel(g)
    obs = make_observations(g, nz, nstars or default_stars[name], seed=seed + 1)
    return g, nz, obs
