"""
TEST INFRASTRUCTURE -- recipe for ``oracle/_ref``: the reference's own Python files on the
fitting path, taken from /root/reference where they lie, so that the GPU box (which has no
/root/reference) can time the *reference itself* as the CPU arm of ``bench.py``.

    python -m oracle.make_ref          (``__graft_entry__.build()`` runs it when the tree is there)

``oracle/_ref/`` is git-ignored (no reference source enters the history) but travels with the
snapshot.  The files are pure Python (the reference has no compiled code);
they run under the import stand-ins of ``oracle/ref_shim.py`` exactly like the tree they come
from.  Nothing in the product imports them.
"""
import os
import shutil
import sys

SRC = os.environ.get("BEAST_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")

FILES = [
    "beast/__init__.py",
    "beast/fitting/__init__.py",
    "beast/fitting/fit.py",
    "beast/fitting/pdf1d.py",
    "beast/fitting/pdf2d.py",
    "beast/fitting/trim_grid.py",
    "beast/fitting/fit_metrics/__init__.py",
    "beast/fitting/fit_metrics/likelihood.py",
    "beast/fitting/fit_metrics/common.py",
    "beast/tools/__init__.py",
    "beast/tools/symlog.py",
    "beast/tools/subgridding_tools.py",
    "beast/tools/merge_beast_stats.py",
    "beast/physicsmodel/__init__.py",
    "beast/observationmodel/__init__.py",
    "beast/observationmodel/noisemodel/__init__.py",
]


def main():
    if not os.path.isfile(os.path.join(SRC, "beast", "fitting", "fit.py")):
        print("oracle/make_ref: no reference tree at %s, nothing done" % SRC)
        return 1
    for rel in FILES:
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), dst)
    print("oracle/make_ref: %d files -> %s" % (len(FILES), DST))
    return 0


if __name__ == "__main__":
    sys.exit(main())
