"""The C-ABI library loads and exports every symbol include/beast_b200.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

import pytest

from tests.util import ROOT


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "beast_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(bf_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    assert "bf_dense_chi2" in syms and "bf_exact_accumulate" in syms and len(syms) >= 15


def test_library_exports_every_declared_symbol():
    from beast_b200 import _cabi

    if not os.path.isfile(_cabi.LIB_PATH):
        import __graft_entry__ as ge

        ge.build()
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), "missing export: " + name


def test_prototypes_cover_header():
    from beast_b200 import _cabi

    assert sorted(_cabi.PROTOTYPES) == declared_symbols()
    lib = _cabi.load()
    assert lib.bf_version() >= 100
    assert _cabi.TILE_MODELS == 256 and _cabi.GROUP == 64 and _cabi.CHUNK == 8


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from beast_b200 import _cabi

    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _cabi.load()


def test_engine_refuses_to_run_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from beast_b200 import synth
    from beast_b200.engine import FitEngine
    from beast_b200.gridprep import FitSetup

    g = synth.make_sedgrid("tiny", seed=1)
    nz = synth.make_noisemodel(g)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FitEngine(FitSetup(g, nz, ["Av", "M_ini"], max_nbins=20))
