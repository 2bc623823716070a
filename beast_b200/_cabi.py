"""
ctypes binding of ``libbeast_b200.so`` (declarations: ``include/beast_b200.h``).

The library is the product; there is no CPU fallback.  ``load()`` raises
``RuntimeError`` when the shared object is missing so that a GPU box without the
built extension fails loudly instead of silently computing something else.
Importing this module does not touch CUDA (fork-safe, like the reference's
``multiprocessing`` callers need -- SURVEY.md section 8b).
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbeast_b200.so")

TILE_MODELS = 256
GROUP = 64
CHUNK = 8
STAR_BLOCK = 1024
MAX_BANDS = 16
MAX_PARAMS = 128
SELECT_SPLIT = 16

c_void_p = C.c_void_p
i64 = C.c_int64


class ExactGrid(C.Structure):
    _fields_ = [
        ("nb", C.c_int),
        ("n_models", i64),
        ("ld_model", i64),
        ("msb", c_void_p),
        ("ivar", c_void_p),
        ("lnq", c_void_p),
        ("lnw", c_void_p),
        ("cov_off", c_void_p),
    ]


class ParamTables(C.Structure):
    _fields_ = [
        ("n_q", C.c_int),
        ("total_bins", C.c_int),
        ("max_bins", C.c_int),
        ("bin_tab", c_void_p),
        ("q_tab", c_void_p),
        ("q_lin", c_void_p),
        ("bin_off", c_void_p),
        ("bin_val", c_void_p),
        ("q_const", c_void_p),
        ("bin_q", c_void_p),
        ("n_fast", C.c_int),
        ("n_slow", C.c_int),
        ("fast_pad", C.c_int),
        ("slow_pad", C.c_int),
        ("fast_k", c_void_p),
        ("slow_k", c_void_p),
        ("fast_bin", c_void_p),
        ("fast_q", c_void_p),
        ("slow_bin", c_void_p),
        ("slow_q", c_void_p),
        ("slow_run", c_void_p),
        ("fast_rng", c_void_p),
    ]


class AccumScratch(C.Structure):
    _fields_ = [
        ("weights", c_void_p),
        ("part_wsum", c_void_p),
        ("part_chi2", c_void_p),
        ("part_chi2_j", c_void_p),
        ("star_flags", c_void_p),
        ("group_order", c_void_p),
        ("chunk_refs", c_void_p),
        ("chunk_wsum", c_void_p),
        ("exp_part", c_void_p),
    ]


class StarOutputs(C.Structure):
    _fields_ = [
        ("total_log_norm", c_void_p),
        ("chi2min", c_void_p),
        ("chi2min_j", c_void_p),
        ("n_sparse", c_void_p),
        ("exp_vals", c_void_p),
        ("pdf1d", c_void_p),
        ("per_vals", c_void_p),
        ("member_mask", c_void_p),
        ("member_prefix", c_void_p),
    ]


class FitPlan(C.Structure):
    _fields_ = [
        ("nb", C.c_int),
        ("full_cov", C.c_int),
        ("n_groups", i64),
        ("n_tiles", i64),
        ("tiles", c_void_p),
        ("n_tiles_coarse", i64),
        ("tiles_coarse", c_void_p),
        ("inv_band_scale", c_void_p),
        ("err_a", C.c_float),
        ("err_b", C.c_float),
        ("grid", ExactGrid),
        ("params", ParamTables),
        ("n_pairs", C.c_int),
        ("pair_bin", c_void_p),
        ("pair_off", c_void_p),
        ("pair_len", i64),
        ("sm_count", C.c_int),
        ("force_legacy", C.c_int),
        ("max_range_entries", i64),
        ("boxes", c_void_p),
        ("sort_band", C.c_int),
    ]


class FitOutputsC(C.Structure):
    _fields_ = [(n, c_void_p) for n in (
        "pmax", "pmax_j", "total_log_norm", "chi2min", "chi2min_j", "n_sparse", "n_groups_selected",
        "exp_vals", "pdf1d", "per_vals", "pdf2d", "sparse_off", "sparse_j", "sparse_lnp",
        "sparse_chi2")] + [("sparse_capacity", i64)]


class FitState(C.Structure):
    _fields_ = [
        ("flux64", c_void_p),
        ("offsets", c_void_p),
        ("list", c_void_p),
        ("member_mask", c_void_p),
        ("member_prefix", c_void_p),
        ("perm", c_void_p),
        ("n_sparse_sorted", c_void_p),
        ("n_list", i64),
        ("n_records", i64),
        ("launches", i64),
        ("split", C.c_int),
        ("n_ranges", C.c_int),
        ("dense_evaluated", i64),
        ("want_timing", C.c_int),
        ("stage_ms", C.c_float * 12),
    ]


STAGES = ("presort", "dense", "select", "exact_max", "group_order", "exact_weights", "histograms",
          "expectations", "pdf2d", "reorder", "sparse_records")

# name -> (restype, argtypes); every symbol declared in include/beast_b200.h
PROTOTYPES = {
    "bf_last_error": (C.c_char_p, []),
    "bf_version": (C.c_int, []),
    "bf_device_info": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "bf_dense_chi2": (C.c_int, [C.c_int, i64, c_void_p, i64, i64, c_void_p, c_void_p, c_void_p,
                                C.c_int, c_void_p]),
    "bf_dense_chi2_pruned": (C.c_int, [C.c_int, i64, c_void_p, c_void_p, i64, i64, c_void_p, c_void_p, c_void_p,
                                       c_void_p, C.c_int, c_void_p, c_void_p]),
    "bf_dense_chi2_cov": (C.c_int, [C.c_int, i64, c_void_p, i64, i64, c_void_p, c_void_p, c_void_p,
                                    C.c_int, c_void_p]),
    "bf_dense_tiles_per_item": (C.c_int, [i64, i64, C.c_int]),
    "bf_select_count": (C.c_int, [i64, i64, i64, c_void_p, c_void_p, C.c_float, C.c_float,
                                  C.c_float, c_void_p, c_void_p]),
    "bf_select_fill": (C.c_int, [i64, i64, i64, c_void_p, c_void_p, C.c_float, C.c_float,
                                 C.c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_scan_scratch_elems": (i64, [i64]),
    "bf_exclusive_scan_u32": (C.c_int, [i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_exclusive_scan_u32_noalloc": (C.c_int, [i64, c_void_p, c_void_p, c_void_p]),
    "bf_group_order_work_elems": (i64, [i64]),
    "bf_group_order": (C.c_int, [i64, c_void_p, i64, i64, c_void_p, c_void_p, c_void_p]),
    "bf_fit_workspace_bytes": (i64, [C.POINTER(FitPlan), i64, i64]),
    "bf_fit_batch": (C.c_int, [C.POINTER(FitPlan), i64, c_void_p, C.c_int, C.c_double,
                               C.POINTER(C.c_double), C.c_int, i64, C.c_uint64, i64, c_void_p,
                               C.c_size_t, C.POINTER(FitOutputsC), C.POINTER(FitState), c_void_p,
                               c_void_p]),
    "bf_fit_sparse_records": (C.c_int, [C.POINTER(FitPlan), C.POINTER(FitState), i64, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_exact_max": (C.c_int, [C.POINTER(ExactGrid), i64, i64, c_void_p, c_void_p, c_void_p,
                               C.c_float, C.c_float, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p]),
    "bf_exact_accumulate": (C.c_int, [C.POINTER(ExactGrid), C.POINTER(ParamTables), i64, i64, c_void_p,
                                      c_void_p, C.c_double, c_void_p, c_void_p, c_void_p, c_void_p, i64,
                                      i64,
                                      C.POINTER(C.c_double), C.POINTER(AccumScratch),
                                      C.POINTER(StarOutputs), c_void_p, c_void_p]),
    "bf_sparse_gather": (C.c_int, [C.POINTER(ExactGrid), i64, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "bf_sparse_sample_ranks": (C.c_int, [i64, c_void_p, c_void_p, C.c_uint64, i64, c_void_p,
                                         c_void_p]),
    "bf_pdf2d_from_weights": (C.c_int, [i64, C.c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, i64, i64, c_void_p, c_void_p]),
    "bf_pdf2d_accumulate": (C.c_int, [C.POINTER(ExactGrid), C.c_int, c_void_p, c_void_p, i64,
                                      c_void_p, c_void_p, c_void_p, C.c_double, c_void_p,
                                      c_void_p, c_void_p, c_void_p]),
    "bf_trim_flags": (C.c_int, [C.c_int, i64, c_void_p, c_void_p, c_void_p, c_void_p, C.c_double,
                                c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_trim_select": (C.c_int, [i64, c_void_p, C.c_uint32, c_void_p, c_void_p]),
    "bf_trim_compact": (C.c_int, [i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_merge_weights": (C.c_int, [C.c_int, i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bf_scale_rows": (C.c_int, [i64, i64, c_void_p, c_void_p, c_void_p]),
    "bf_axpy_rows": (C.c_int, [i64, i64, c_void_p, c_void_p, c_void_p, C.c_int, c_void_p]),
    "bf_merge_finalize": (C.c_int, [C.POINTER(ParamTables), i64, c_void_p, C.POINTER(C.c_double),
                                    c_void_p, c_void_p]),
}

_lib = None


class BeastB200Error(RuntimeError):
    pass


def load():
    """dlopen the extension once and attach prototypes; raise if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            "beast_b200: CUDA extension %s is missing -- build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` or `make -C beast_b200/csrc`. "
            "There is no CPU fallback." % LIB_PATH
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here = header / library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().bf_last_error().decode("utf-8", "replace")
        raise BeastB200Error("%s failed (%d): %s" % (what or "beast_b200 call", status, msg))


def call(name, *args):
    """invoke an int-returning entry point and raise on a non-zero status"""
    check(getattr(load(), name)(*args), name)
