"""
Dependency-free FITS writer/reader for the three FITS products of the fit
(reference writers: ``beast/fitting/fit.py:40-185``).

astropy is not available on the target image, so the small subset of FITS the
reference emits is produced by hand:

* stats file  : empty primary HDU, BINTABLE HDU1 (one row per star), BINTABLE HDU2
                (``filternames``, ``wavelengths``)                      fit.py:110-125
* pdf1d/pdf2d : primary HDU holding a 2x2 zero image, then one IMAGE extension per
                parameter / parameter pair, ``EXTNAME`` = name           fit.py:147-185

Consumers in the reference read ``ext.data[:-1]`` as stars and ``ext.data[-1]`` as the
bin centres (``tools/star_type_probability.py:77-79``); array shapes are kept.
"""

from __future__ import annotations

import numpy as np

BLOCK = 2880


def _card(key, value=None, comment=""):
    """one or more 80-character header cards.  A string that does not fit the 68 characters a
    card offers is continued with the long-string convention (FITS 4.0 section 4.2.1.2: '&' at
    the end of the value, CONTINUE cards after it) instead of being cut."""
    if value is None:
        s = key.ljust(8)
    elif isinstance(value, bool):
        s = "%-8s= %20s" % (key, "T" if value else "F")
    elif isinstance(value, (int, np.integer)):
        s = "%-8s= %20d" % (key, int(value))
    elif isinstance(value, (float, np.floating)):
        s = "%-8s= %20s" % (key, repr(float(value)).upper())
    else:
        v = str(value).replace("'", "''")
        if len(v) <= 68:
            s = "%-8s= '%-8s'" % (key, v)
        else:
            pieces, rest = [], v
            while len(rest) > 67:
                cut = 67
                if rest[cut - 1] == "'" and (len(rest[:cut]) - len(rest[:cut].rstrip("'"))) % 2 == 1:
                    cut -= 1  # never split a doubled quote
                pieces.append(rest[:cut])
                rest = rest[cut:]
            pieces.append(rest)
            cards = ["%-8s= '%s&'" % (key, pieces[0])]
            for i, pc in enumerate(pieces[1:], start=1):
                cards.append("CONTINUE  '%s%s'" % (pc, "&" if i < len(pieces) - 1 else ""))
            return "".join(c.ljust(80) for c in cards)
    if comment and len(s) + 3 + len(comment) <= 80:
        s += " / " + comment
    if len(s) > 80:
        raise ValueError("FITS header card too long: %r" % s)
    return s.ljust(80)


def _header_bytes(cards):
    txt = "".join(cards) + "END".ljust(80)
    pad = (-len(txt)) % BLOCK
    return (txt + " " * pad).encode("ascii")


def _pad_data(raw):
    pad = (-len(raw)) % BLOCK
    return raw + b"\0" * pad


def _image_cards(arr, first):
    arr = np.asarray(arr)
    bitpix = {"f8": -64, "f4": -32, "i8": 64, "i4": 32, "i2": 16, "u1": 8}[arr.dtype.str[1:]]
    cards = [_card("SIMPLE", True, "conforms to FITS standard")] if first else [
        _card("XTENSION", "IMAGE", "Image extension")]
    cards += [_card("BITPIX", bitpix), _card("NAXIS", arr.ndim)]
    for i, n in enumerate(arr.shape[::-1]):
        cards.append(_card("NAXIS%d" % (i + 1), n))
    if first:
        cards.append(_card("EXTEND", True))
    else:
        cards += [_card("PCOUNT", 0), _card("GCOUNT", 1)]
    return cards


def write_image_file(path, arrays, names):
    """primary 2x2 zero image + one IMAGE extension per array (save_pdf1d / save_pdf2d)"""
    with open(path, "wb") as f:
        prim = np.zeros((2, 2))
        f.write(_header_bytes(_image_cards(prim, True)))
        f.write(_pad_data(prim.astype(">f8").tobytes()))
        for arr, name in zip(arrays, names):
            arr = np.asarray(arr, dtype=np.float64)
            cards = _image_cards(arr, False) + [_card("EXTNAME", name, "extension name")]
            f.write(_header_bytes(cards))
            f.write(_pad_data(arr.astype(">f8").tobytes()))


def _table_hdu(columns, extname=None, extra=None):
    """columns: ordered dict name -> 1-D array (float / int / str); extra: header keywords"""
    names = list(columns.keys())
    nrows = len(next(iter(columns.values()))) if names else 0
    fields = []
    for n in names:
        a = np.asarray(columns[n])
        if a.dtype.kind in "US":
            width = max(1, max((len(str(x)) for x in a), default=1))
            fields.append((n, "%dA" % width, "S%d" % width, np.array([str(x) for x in a], dtype="S%d" % width)))
        elif a.dtype.kind in "iu":
            fields.append((n, "K", ">i8", a.astype(">i8")))
        elif a.dtype.kind == "b":
            fields.append((n, "L", "S1", np.where(a, b"T", b"F").astype("S1")))
        else:
            fields.append((n, "D", ">f8", a.astype(">f8")))
    dt = np.dtype([("f%d" % i, f[2]) for i, f in enumerate(fields)])
    rec = np.zeros(nrows, dtype=dt)
    for i, f in enumerate(fields):
        rec["f%d" % i] = f[3]
    cards = [_card("XTENSION", "BINTABLE", "binary table extension"), _card("BITPIX", 8),
             _card("NAXIS", 2), _card("NAXIS1", dt.itemsize), _card("NAXIS2", nrows),
             _card("PCOUNT", 0), _card("GCOUNT", 1), _card("TFIELDS", len(fields))]
    for i, f in enumerate(fields):
        cards.append(_card("TTYPE%d" % (i + 1), f[0]))
        cards.append(_card("TFORM%d" % (i + 1), f[1]))
    if extname is not None:
        cards.append(_card("EXTNAME", extname, "extension name"))
    for k, v in (extra or {}).items():
        cards.append(_card(k.upper()[:8], v))
    return _header_bytes(cards) + _pad_data(rec.tobytes())


def write_table_file(path, tables):
    """empty primary HDU + one BINTABLE per dict in ``tables`` (save_stats)"""
    with open(path, "wb") as f:
        f.write(_header_bytes([_card("SIMPLE", True, "conforms to FITS standard"), _card("BITPIX", 8),
                               _card("NAXIS", 0), _card("EXTEND", True)]))
        for t in tables:
            f.write(_table_hdu(t))


# ------------------------------------------------------------------------------
def _quoted(v):
    """text of a quoted FITS string whose opening quote has been consumed"""
    out, j = "", 0
    while j < len(v):
        if v[j] == "'":
            if j + 1 < len(v) and v[j + 1] == "'":
                out += "'"
                j += 2
                continue
            break
        out += v[j]
        j += 1
    return out.rstrip()


def _parse_header(buf, pos):
    cards = {}
    order = []
    while True:
        block = buf[pos:pos + BLOCK].decode("ascii")
        pos += BLOCK
        done = False
        for i in range(0, BLOCK, 80):
            c = block[i:i + 80]
            key = c[:8].strip()
            if key == "END":
                done = True
                break
            if key == "CONTINUE" and order and isinstance(cards.get(order[-1]), str):
                prev = cards[order[-1]]
                if prev.endswith("&"):  # long-string convention
                    v = c[8:].lstrip()
                    piece = _quoted(v[1:]) if v.startswith("'") else ""
                    cards[order[-1]] = prev[:-1] + piece
                continue
            if c[8:10] != "= ":
                continue
            val = c[10:]
            if val.lstrip().startswith("'"):
                cards[key] = _quoted(val.lstrip()[1:])
            else:
                v = val.split("/")[0].strip()
                if v in ("T", "F"):
                    cards[key] = v == "T"
                else:
                    try:
                        cards[key] = int(v)
                    except ValueError:
                        cards[key] = float(v)
            order.append(key)
        if done:
            return cards, pos


def read_fits(path):
    """list of (header dict, data) for every HDU; tables come back as dict name -> array"""
    with open(path, "rb") as f:
        buf = f.read()
    pos, hdus = 0, []
    while pos < len(buf):
        hdr, pos = _parse_header(buf, pos)
        naxis = hdr.get("NAXIS", 0)
        shape = [hdr["NAXIS%d" % (i + 1)] for i in range(naxis)][::-1]
        nbytes = abs(hdr["BITPIX"]) // 8 * int(np.prod(shape)) if naxis else 0
        raw = buf[pos:pos + nbytes]
        pos += nbytes + ((-nbytes) % BLOCK)
        if hdr.get("XTENSION", "IMAGE").strip() == "BINTABLE":
            nf = hdr["TFIELDS"]
            kinds = {"D": ">f8", "K": ">i8", "E": ">f4", "J": ">i4", "L": "S1"}
            dts = []
            for i in range(nf):
                tf = hdr["TFORM%d" % (i + 1)].strip()
                dts.append(("f%d" % i, "S%d" % int(tf[:-1] or 1) if tf.endswith("A") else kinds[tf[-1]]))
            rec = np.frombuffer(raw, dtype=np.dtype(dts), count=hdr["NAXIS2"])
            data = {}
            for i in range(nf):
                col = rec["f%d" % i]
                if col.dtype.kind == "S":
                    col = np.array([x.decode("ascii").rstrip() for x in col])
                else:
                    col = col.astype(col.dtype.newbyteorder("="))
                data[hdr["TTYPE%d" % (i + 1)].strip()] = col
        elif naxis:
            dt = {-64: ">f8", -32: ">f4", 64: ">i8", 32: ">i4", 16: ">i2", 8: "u1"}[hdr["BITPIX"]]
            data = np.frombuffer(raw, dtype=dt).reshape(shape).astype(np.dtype(dt).newbyteorder("="))
        else:
            data = None
        hdus.append((hdr, data))
    return hdus


# ------------------------------------------------------------------------------
# SED grid files in the reference's FITS layout (physicsmodel/helpers/gridbackends.py:151-177
# write, 345-365 read): primary HDU = wavelengths, IMAGE "seds" (n_models x n_bands), optional
# IMAGE "covdiag" / "covoffdiag", BINTABLE "grid" with one column per model property and the
# filter names in its header keyword FILTERS (space separated).
class FitsSEDGrid:
    """what the fit touches of a SEDGrid (fit.py:396-463), read from a FITS grid file"""

    def __init__(self, lamb, seds, cols, filters, cov_diag=None, cov_offdiag=None, header=None):
        self.lamb = np.asarray(lamb, dtype=float)
        self.seds = np.asarray(seds, dtype=float)
        self._cols = dict(cols)
        self.filters = list(filters) if filters is not None else None
        self.cov_diag, self.cov_offdiag = cov_diag, cov_offdiag
        self.header = dict(header or {})

    def keys(self):
        return list(self._cols.keys())

    def __getitem__(self, name):
        return self._cols[name]

    def __contains__(self, name):
        return name in self._cols

    def __len__(self):
        return self.seds.shape[0]


def read_sedgrid_fits(path):
    """read a SED grid written by the reference's ``writeFITS`` (or by ``write_sedgrid_fits``)"""
    hdus = read_fits(path)
    by_name = {str(h.get("EXTNAME", "")).strip().lower(): (h, d) for h, d in hdus[1:]}
    if "seds" not in by_name or "grid" not in by_name:
        raise ValueError("%s is not a SED grid in the seds/grid FITS layout" % path)
    ghdr, gtab = by_name["grid"]
    filters = str(ghdr["FILTERS"]).split() if "FILTERS" in ghdr else None
    return FitsSEDGrid(hdus[0][1], by_name["seds"][1], gtab, filters,
                       by_name.get("covdiag", (None, None))[1], by_name.get("covoffdiag", (None, None))[1],
                       {k: v for k, v in ghdr.items()})


def write_sedgrid_fits(path, grid):
    """write ``grid`` (lamb, seds, keys()/[name], filters) in the same layout"""
    with open(path, "wb") as f:
        lamb = np.asarray(grid.lamb, dtype=np.float64)
        f.write(_header_bytes(_image_cards(lamb, True)))
        f.write(_pad_data(lamb.astype(">f8").tobytes()))
        images = [("seds", grid.seds)]
        for name in ("cov_diag", "cov_offdiag"):
            if getattr(grid, name, None) is not None:
                images.append((name.replace("_", ""), getattr(grid, name)))
        for name, arr in images:
            arr = np.asarray(arr, dtype=np.float64)
            f.write(_header_bytes(_image_cards(arr, False) + [_card("EXTNAME", name, "extension name")]))
            f.write(_pad_data(arr.astype(">f8").tobytes()))
        extra = {"FILTERS": " ".join(grid.filters)} if getattr(grid, "filters", None) else None
        f.write(_table_hdu({k: grid[k] for k in grid.keys()}, extname="grid", extra=extra))
