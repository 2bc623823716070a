"""The FITS files this package writes, checked against the FITS 4.0 standard by a validator
written from the standard's text -- independent of the package's own reader: block structure,
card syntax, mandatory keywords and their order, table column arithmetic, long strings."""
import os

import numpy as np
import pytest

BLOCK = 2880


def cards_of(buf, pos):
    """(list of 80-character cards up to END, position after the header blocks)"""
    cards = []
    while True:
        block = buf[pos:pos + BLOCK]
        assert len(block) == BLOCK, "header runs past the end of the file"
        pos += BLOCK
        text = block.decode("ascii")  # section 4.1: 7-bit ASCII
        assert all(32 <= ord(ch) <= 126 for ch in text), "non-printable character in a header"
        for i in range(0, BLOCK, 80):
            c = text[i:i + 80]
            cards.append(c)
            if c[:8] == "END     ":
                assert c[8:].strip() == ""
                assert text[i + 80:].strip() == "", "header not padded with blanks after END"
                return cards[:-1], pos


def value_of(card):
    """parsed value of a keyword card (section 4.2) -- strings, logicals, integers, floats"""
    assert card[8:10] == "= ", card
    body = card[10:]
    if body.lstrip().startswith("'"):
        start = body.index("'")
        assert start == 0, "the opening quote of a string sits in column 11"
        j, out = 1, ""
        while True:
            assert j < len(body), "unterminated string: " + card
            if body[j] == "'":
                if j + 1 < len(body) and body[j + 1] == "'":
                    out += "'"
                    j += 2
                    continue
                break
            out += body[j]
            j += 1
        assert body[j + 1:].strip() == "" or body[j + 1:].lstrip().startswith("/"), card
        return out.rstrip()
    v = body.split("/")[0].strip()
    if v in ("T", "F"):
        assert body[:20].rstrip().endswith(v) and len(body[:20].rstrip()) == 20, "logical in column 30"
        return v == "T"
    assert len(body.split("/")[0].rstrip()) <= 20 or "." in v or "E" in v, card
    return int(v) if v.lstrip("+-").isdigit() else float(v)


def validate(path):
    """walks the HDUs; returns [(keywords dict incl. joined long strings, data bytes)]"""
    buf = open(path, "rb").read()
    assert len(buf) % BLOCK == 0, "file length is not a multiple of 2880"
    pos, hdus = 0, []
    while pos < len(buf):
        cards, pos = cards_of(buf, pos)
        keys, order, last = {}, [], None
        for c in cards:
            k = c[:8].rstrip()
            assert k == k.upper() and all(ch.isalnum() or ch in "-_" for ch in k), "bad keyword " + repr(k)
            if k == "CONTINUE":
                assert last is not None and isinstance(keys[last], str) and keys[last].endswith("&")
                assert c[8:10] == "  " and c[10] == "'"
                piece = value_of(c[:8] + "= " + c[10:])
                keys[last] = keys[last][:-1] + piece
                continue
            if k in ("COMMENT", "HISTORY", ""):
                continue
            keys[k] = value_of(c)
            order.append(k)
            last = k
        for k, v in keys.items():
            assert not (isinstance(v, str) and v.endswith("&")), "dangling long-string marker in " + k
        first = not hdus
        # mandatory keywords in order (tables 7, 10, 17 of the standard)
        assert order[0] == ("SIMPLE" if first else "XTENSION")
        assert order[1] == "BITPIX" and order[2] == "NAXIS"
        naxis = keys["NAXIS"]
        assert order[3:3 + naxis] == ["NAXIS%d" % (i + 1) for i in range(naxis)]
        if first:
            assert keys["SIMPLE"] is True
        else:
            assert order[3 + naxis:5 + naxis] == ["PCOUNT", "GCOUNT"]
            assert keys["GCOUNT"] == 1 and keys["PCOUNT"] == 0
        assert keys["BITPIX"] in (8, 16, 32, 64, -32, -64)
        n = 1
        for i in range(naxis):
            n *= keys["NAXIS%d" % (i + 1)]
        nbytes = abs(keys["BITPIX"]) // 8 * n if naxis else 0
        if not first and keys["XTENSION"] == "BINTABLE":
            assert order[5 + naxis] == "TFIELDS" and keys["BITPIX"] == 8 and naxis == 2
            width = 0
            for i in range(keys["TFIELDS"]):
                tform = keys["TFORM%d" % (i + 1)]
                assert "TTYPE%d" % (i + 1) in keys
                rep = int(tform[:-1] or 1)
                width += rep * {"L": 1, "A": 1, "B": 1, "I": 2, "J": 4, "K": 8, "E": 4, "D": 8}[tform[-1]]
            assert width == keys["NAXIS1"], "TFORMn do not add up to NAXIS1"
        data = buf[pos:pos + nbytes]
        assert len(data) == nbytes
        pad = (-nbytes) % BLOCK
        assert buf[pos + nbytes:pos + nbytes + pad] == b"\0" * pad, "data not zero padded"
        pos += nbytes + pad
        hdus.append((keys, data))
    return hdus


def test_image_and_table_files_are_valid_fits(tmp_path):
    from beast_b200 import fitsio

    long_name = "symlogHST_ACS_WFC_F475W_wd_bias_with_a_very_long_tail_that_does_not_fit_a_single_card_x"
    assert len(long_name) > 68
    p1 = str(tmp_path / "pdf1d.fits")
    fitsio.write_image_file(p1, [np.arange(12.0).reshape(3, 4), np.ones((2, 5))], ["Av", long_name])
    h = validate(p1)
    assert [k.get("EXTNAME") for k, _ in h[1:]] == ["Av", long_name]  # not cut at column 80
    assert np.array_equal(np.frombuffer(h[1][1], dtype=">f8").reshape(3, 4), np.arange(12.0).reshape(3, 4))
    p2 = str(tmp_path / "stats.fits")
    cols = {"Name": np.array(["a", "it's", "ccc"]), long_name + "_Best": np.array([1.0, 2.0, 3.0]),
            "Pmax_indx": np.array([1, 2, 3]), "ok": np.array([True, False, True])}
    fitsio.write_table_file(p2, [cols, {"filternames": np.array(["F1"]), "wavelengths": np.array([1.0])}])
    h = validate(p2)
    assert h[1][0]["TTYPE2"] == long_name + "_Best" and h[1][0]["NAXIS2"] == 3
    # and the package's reader agrees with what the validator saw
    back = fitsio.read_fits(p2)
    assert list(back[1][1].keys()) == list(cols.keys())
    assert list(back[1][1]["Name"]) == ["a", "it's", "ccc"]
    assert fitsio.read_fits(p1)[2][0]["EXTNAME"] == long_name


def test_sedgrid_fits_round_trip(tmp_path):
    """a SED grid in the reference's FITS layout (gridbackends.py:151-177): primary = wavelengths,
    IMAGE seds, BINTABLE grid with the filters in its header"""
    from beast_b200 import fit, fitsio, synth

    g = synth.make_sedgrid("tiny", seed=3)
    path = str(tmp_path / "grid_seds.grid.fits")
    fitsio.write_sedgrid_fits(path, g)
    h = validate(path)
    assert [k.get("EXTNAME") for k, _ in h[1:]] == ["seds", "grid"]
    assert h[0][0]["NAXIS1"] == len(g.lamb) and h[1][0]["NAXIS1"] == g.seds.shape[1]
    assert h[2][0]["FILTERS"] == " ".join(g.filters)  # (longer than one card: continued)
    back = fit._load_grid(path, "memory")
    assert back.filters == g.filters and back.keys() == g.keys()
    np.testing.assert_array_equal(back.seds, g.seds)
    np.testing.assert_array_equal(back.lamb, g.lamb)
    for k in g.keys():
        np.testing.assert_array_equal(back[k], g[k])
