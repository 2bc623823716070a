"""
File-level merges that follow the fit (row f4 of SURVEY.md section 8).

``merge_stats_files`` keeps the signature and behaviour of
``beast.tools.merge_beast_stats.merge_stats_files`` (reference ``beast/tools/merge_beast_stats.py:12-71``):
stack the stats tables of several fits (source-density bins / sub-files), tag every row with the
file it came from (``reorder_tag``), carry the filter table of the first file along, and return
the number of rows.  It uses the package's own FITS reader/writer, so it runs where astropy is
not installed.
"""

from __future__ import annotations

import numpy as np

from . import fitsio


def merge_stats_files(stats_files, out_stats_filebase, reorder_tag_list=None):
    filters_tab = None
    first = fitsio.read_fits(stats_files[0])
    if len(first) > 2 and isinstance(first[2][1], dict):
        filters_tab = first[2][1]
    tables = []
    for i, cur_stat in enumerate(stats_files):
        cur = dict(fitsio.read_fits(cur_stat)[1][1])
        if reorder_tag_list is None:
            # the tag is what sits between "_sd" and "_stats" in the file name (:47-49)
            bpos = cur_stat.find("_sd") + 1
            epos = cur_stat.find("_stats")
            tag = cur_stat[bpos:epos]
        else:
            tag = reorder_tag_list[i]
        n = len(next(iter(cur.values()))) if cur else 0
        cur["reorder_tag"] = np.array([tag] * n)
        tables.append(cur)
    names = list(tables[0].keys())
    for t in tables[1:]:
        if list(t.keys()) != names:
            raise ValueError("stats files to merge do not have the same columns")
    full = {k: np.concatenate([np.asarray(t[k]) for t in tables]) for k in names}
    out = [full] + ([filters_tab] if filters_tab is not None else [])
    fitsio.write_table_file(out_stats_filebase + "_stats.fits", out)
    return len(next(iter(full.values()))) if full else 0
