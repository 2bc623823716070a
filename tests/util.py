"""Shared helpers of the parity tests."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["tiny_thr10_pdf2d", "tiny_thr40_nonorm", "small_thr10", "tiny_fullcov"]

# parity tolerances (BASELINE.json north_star; reference regression test helpers.py:33-112)
LNP_ATOL = 1e-4       # lnP-like quantities: Pmax, total_log_norm, sparse lnp
PDF_RTOL = 1e-5       # pdf1d / pdf2d bins (atol for empty bins)
PDF_ATOL = 1e-12
STAT_RTOL = 1e-6      # expectation / chi2min, the reference's own stats tolerance


def load_golden(name):
    from oracle.make_golden import unpack_inputs

    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    g, nz, obs = unpack_inputs(z)
    ref = {k[4:]: z[k] for k in z.files if k.startswith("ref_") and not k.startswith("ref_pdf")}
    ref["pdf1d"] = [z[k] for k in sorted(f for f in z.files if f.startswith("ref_pdf1d_"))]
    ref["pdf2d"] = [z[k] for k in sorted(f for f in z.files if f.startswith("ref_pdf2d_") and f != "ref_pdf2d_pairs")]
    ref["pdf2d_pairs"] = [str(s) for s in z["ref_pdf2d_pairs"]]
    ref["qnames"] = [str(s) for s in z["ref_qnames"]]
    args = dict(
        keys=[str(s) for s in z["keys"]],
        threshold=float(z["threshold"]),
        max_nbins=int(z["max_nbins"]),
        do_not_normalize=bool(z["do_not_normalize"]),
        pdf2d_param_list=[str(s) for s in z["pdf2d_param_list"]] or None,
    )
    return g, nz, obs, args, ref


def bin_width_tolerance(bin_vals):
    """one bin width (the north-star tolerance for percentiles)"""
    if len(bin_vals) < 2:
        return 1e-9 + 1e-9 * abs(float(bin_vals[0]))
    return float(np.max(np.abs(np.diff(bin_vals))))


def compare_to_reference(setup, res, ref, n_pdf2d=0):
    """CUDA result (engine.FitResult) against reference-shaped arrays (make_golden / fit_all)."""
    g0 = setup.g0_indxs
    assert list(setup.qnames) == list(ref["qnames"])
    # best-fit indices: bit exact
    np.testing.assert_array_equal(g0[res.Pmax_j], ref["Pmax_indx"].astype(np.int64))
    np.testing.assert_array_equal(g0[res.chi2min_j], ref["chi2min_indx"].astype(np.int64))
    np.testing.assert_allclose(res.Pmax, ref["Pmax"], rtol=0, atol=LNP_ATOL)
    np.testing.assert_allclose(res.total_log_norm, ref["total_log_norm"], rtol=0, atol=LNP_ATOL)
    np.testing.assert_allclose(res.chi2min, ref["chi2min"], rtol=STAT_RTOL, atol=1e-9)
    # expectations: FP64 sums over 32-bit fixed-point parameter values (2^-33 of the range)
    span = np.array([c[g0].max() - c[g0].min() for c in setup.qcols])
    assert np.all(np.abs(res.exp_vals - ref["exp_vals"]) <= STAT_RTOL * np.abs(ref["exp_vals"]) + 1e-9 * span + 1e-300)
    best = np.stack([setup.qcols[k][g0[res.Pmax_j]] for k in range(setup.n_q)], axis=1)
    np.testing.assert_array_equal(best, ref["best_vals"])
    np.testing.assert_array_equal(setup.specgrid_indx[g0[res.Pmax_j]], ref["specgrid_indx"].astype(np.int64))
    for k in range(setup.n_q):
        a, b = setup.bin_off[k], setup.bin_off[k + 1]
        np.testing.assert_array_equal(setup.bin_val[a:b], ref["pdf1d"][k][-1])
        np.testing.assert_allclose(res.pdf1d[:, a:b], ref["pdf1d"][k][:-1], rtol=PDF_RTOL, atol=PDF_ATOL,
                                   err_msg="pdf1d of " + setup.qnames[k])
        tol = bin_width_tolerance(setup.bin_val[a:b])
        assert np.max(np.abs(res.per_vals[:, k, :] - ref["per_vals"][:, k, :])) <= tol, setup.qnames[k]
        # in practice the percentiles agree far better than one bin
        np.testing.assert_allclose(res.per_vals[:, k, :], ref["per_vals"][:, k, :], rtol=1e-7,
                                   atol=1e-7 * tol + 1e-300)
    if n_pdf2d:
        for p, pair in enumerate(setup.pairs):
            a, b = setup.pair_off[p], setup.pair_off[p + 1]
            got = res.pdf2d[:, a:b].reshape(-1, pair.nb1, pair.nb2)
            np.testing.assert_allclose(got, ref["pdf2d"][p][:-2], rtol=PDF_RTOL, atol=PDF_ATOL)
            np.testing.assert_array_equal(np.tile(pair.bin_vals_p1, (pair.nb2, 1)).T, ref["pdf2d"][p][-2])
            np.testing.assert_array_equal(np.tile(pair.bin_vals_p2, (pair.nb1, 1)), ref["pdf2d"][p][-1])


def compare_sparse(setup, res, off, idx, lnp, chi2):
    """full sparse sets: indices exact, lnp within LNP_ATOL (stored float32 like the reference)"""
    g0 = setup.g0_indxs
    np.testing.assert_array_equal(res.sparse_off, off)
    np.testing.assert_array_equal(g0[res.sparse_j], idx)
    np.testing.assert_allclose(res.sparse_lnp, lnp.astype(np.float32), rtol=0, atol=LNP_ATOL)
    np.testing.assert_allclose(res.sparse_chi2, chi2.astype(np.float32), rtol=1e-6, atol=1e-6)


class NumpyMergeMath:
    """test double for beast_b200.subgrid.CudaMergeMath (the product's own hooks are CUDA
    kernels): the same four hooks in NumPy, percentiles with the oracle's restated function"""

    def __init__(self, bin_rows, percentiles=(16.0, 50.0, 84.0)):
        self.bin_rows = [np.asarray(b, dtype=float) for b in bin_rows]
        self.set_percentiles(percentiles)

    def set_percentiles(self, percentiles):
        self.pct = tuple(float(x) for x in percentiles)

    def weights(self, tln_all):
        import torch

        t = tln_all.numpy()
        mx = t.max(axis=0)
        w = np.exp(t - mx[None, :])
        return torch.from_numpy(w), torch.from_numpy(np.log(w.sum(axis=0)) + mx)

    def scale_rows(self, rows, w):
        return rows * w.reshape(-1, *([1] * (rows.dim() - 1)))

    def axpy_rows(self, w, src, dst, accumulate):
        v = src * w.reshape(-1, *([1] * (src.dim() - 1)))
        if accumulate:
            dst += v
        else:
            dst.copy_(v)

    def finalize(self, pdf):
        import torch

        from oracle import fit_oracle as fo

        p = pdf.numpy()
        per = np.zeros((p.shape[0], len(self.bin_rows), 3))
        off = 0
        for k, b in enumerate(self.bin_rows):
            blk = p[:, off:off + len(b)]
            norm = blk.sum(axis=1, keepdims=True)
            np.divide(blk, norm, out=blk, where=norm > 0)
            for e in range(p.shape[0]):
                if blk[e].max() > 0:
                    per[e, k] = fo.percentile(b, list(self.pct), blk[e])
            off += len(b)
        return torch.from_numpy(per)
