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
        s = "%-8s= '%-8s'" % (key, v)
    if comment:
        s += " / " + comment
    return s[:80].ljust(80)


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


def _table_hdu(columns):
    """columns: ordered dict name -> 1-D array (float / int / str)"""
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
    return _header_bytes(cards) + _pad_data(rec.tobytes())


def write_table_file(path, tables):
    """empty primary HDU + one BINTABLE per dict in ``tables`` (save_stats)"""
    with open(path, "wb") as f:
        f.write(_header_bytes([_card("SIMPLE", True, "conforms to FITS standard"), _card("BITPIX", 8),
                               _card("NAXIS", 0), _card("EXTEND", True)]))
        for t in tables:
            f.write(_table_hdu(t))


# ------------------------------------------------------------------------------
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
            if c[8:10] != "= ":
                continue
            val = c[10:]
            if val.lstrip().startswith("'"):
                v = val.lstrip()[1:]
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
                cards[key] = out.rstrip()
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
