"""Minimal pure-numpy FITS reader/writer for the subtraction path.

The reference moves every array across the hotpants process boundary as FITS
files (zuds/subtraction.py:68-86,162; hotpants/main.c:120-203,224-373).  astropy
and cfitsio are not available here, so this module implements just what that
boundary needs: 2-D image HDUs (BITPIX 8/16/32/-32/-64 with BSCALE/BZERO),
ordered header cards, CONTINUE long strings and the all-double binary table
hotpants appends with ``-hki`` (main.c:573-579).
"""
from __future__ import annotations

import numpy as np

BLOCK = 2880
_DT = {8: ">u1", 16: ">i2", 32: ">i4", -32: ">f4", -64: ">f8"}
_BITPIX = {"uint8": 8, "int16": 16, "uint16": 16, "int32": 32, "float32": -32, "float64": -64}


class Header:
    """Ordered list of (key, value, comment) cards; dict-style access by key."""

    def remove(self, key):
        self.cards = [c for c in self.cards if c[0] != key]

    def __init__(self, cards=None):
        self.cards = list(cards) if cards else []

    def __contains__(self, key):
        return any(k == key for k, _, _ in self.cards)

    def __getitem__(self, key):
        for k, v, _ in self.cards:
            if k == key:
                return v
        raise KeyError(key)

    def get(self, key, default=None):
        return self[key] if key in self else default

    def __setitem__(self, key, value):
        comment = ""
        if isinstance(value, tuple):
            value, comment = value
        for i, (k, _, c) in enumerate(self.cards):
            if k == key and key not in ("COMMENT", "HISTORY", ""):
                self.cards[i] = (k, value, comment or c)
                return
        self.cards.append((key, value, comment))

    def append(self, key, value, comment=""):
        self.cards.append((key, value, comment))

    def keys(self):
        return [k for k, _, _ in self.cards]

    def copy(self):
        return Header(self.cards)


class Raw(str):
    """A pre-formatted numeric value field (written right-justified, unquoted), for cards whose
    number format is dictated by the reference (cfitsio's %G / %E key writers)."""


def _fmt_value(v):
    if isinstance(v, Raw):
        return "%20s" % str(v)
    if isinstance(v, (bool, np.bool_)):
        return "%20s" % ("T" if v else "F")
    if isinstance(v, (int, np.integer)):
        return "%20d" % int(v)
    if isinstance(v, (float, np.floating)):
        s = repr(float(v)).upper()
        if "." not in s and "E" not in s and "N" not in s and "I" not in s:
            s += "."
        return "%20s" % s
    s = str(v).replace("'", "''")
    return "'%-8s'" % s


def _cards_for(key, value, comment):
    if key in ("COMMENT", "HISTORY", ""):
        return [("%-8s%s" % (key, str(value)))[:80].ljust(80)]
    if isinstance(value, str) and len(value.replace("'", "''")) > 68:
        out, s, first = [], value, True
        while s or first:
            piece, s = s[:67], s[67:]
            q = "'%s%s'" % (piece.replace("'", "''"), "&" if s else "")
            out.append((("%-8s= %s" % (key, q)) if first else ("CONTINUE  %s" % q))[:80].ljust(80))
            first = False
        return out
    card = "%-8s= %s" % (key, _fmt_value(value))
    if comment:
        card += " / " + comment
    return [card[:80].ljust(80)]


def _parse_card(card):
    key = card[:8].rstrip()
    if key in ("COMMENT", "HISTORY", "") or card[8:10] != "= ":
        return key, card[8:].rstrip(), ""
    body = card[10:]
    s = body.lstrip()
    if s.startswith("'"):
        i, val = 1, []
        while i < len(s):
            if s[i] == "'":
                if i + 1 < len(s) and s[i + 1] == "'":
                    val.append("'")
                    i += 2
                    continue
                break
            val.append(s[i])
            i += 1
        rest = s[i + 1:]
        comment = rest.split("/", 1)[1].strip() if "/" in rest else ""
        return key, "".join(val).rstrip(), comment
    val, _, comment = body.partition("/")
    val = val.strip()
    if val == "T":
        return key, True, comment.strip()
    if val == "F":
        return key, False, comment.strip()
    try:
        return key, int(val), comment.strip()
    except ValueError:
        pass
    try:
        return key, float(val.replace("D", "E").replace("d", "e")), comment.strip()
    except ValueError:
        return key, val, comment.strip()


def _read_header(buf, off):
    cards = []
    while True:
        block = buf[off:off + BLOCK]
        if len(block) < BLOCK:
            raise IOError("truncated FITS header")
        off += BLOCK
        text = bytes(block).decode("ascii", "replace")
        done = False
        for i in range(36):
            c = text[i * 80:(i + 1) * 80]
            if c.startswith("END     "):
                done = True
                break
            if c.strip():
                cards.append(c)
        if done:
            break
    hdr = Header()
    for c in cards:
        k, v, cm = _parse_card(c)
        if k == "CONTINUE" and hdr.cards and isinstance(hdr.cards[-1][1], str):
            pk, pv, pc = hdr.cards[-1]
            add = c[8:].strip()
            add = add[1:add.rfind("'")] if add.startswith("'") else add
            if pv.endswith("&"):
                pv = pv[:-1]
            hdr.cards[-1] = (pk, pv + add.replace("''", "'"), pc)
            continue
        hdr.append(k, v, cm)
    for i, (k, v, cm) in enumerate(hdr.cards):
        if isinstance(v, str) and v.endswith("&"):
            hdr.cards[i] = (k, v[:-1], cm)
    return hdr, off


def read_hdus(path):
    """Return a list of (Header, ndarray-or-None) for every HDU in ``path``."""
    buf = np.fromfile(path, dtype=np.uint8)
    out, off = [], 0
    while off < len(buf):
        hdr, off = _read_header(buf, off)
        naxis = int(hdr.get("NAXIS", 0))
        shape = [int(hdr["NAXIS%d" % (i + 1)]) for i in range(naxis)]
        bitpix = int(hdr.get("BITPIX", 8))
        nbytes = (abs(bitpix) // 8) * int(np.prod(shape)) if naxis else 0
        data = None
        if nbytes:
            raw = buf[off:off + nbytes]
            if hdr.get("XTENSION", "").startswith("BINTABLE"):
                nf = int(hdr["TFIELDS"])
                data = raw.view(">f8").reshape(shape[1], nf).astype(np.float64)
            else:
                data = raw.view(_DT[bitpix]).reshape(shape[::-1])
                bs, bz = float(hdr.get("BSCALE", 1.0)), float(hdr.get("BZERO", 0.0))
                if bs != 1.0 or bz != 0.0:
                    data = data.astype(np.float64) * bs + bz
                    if bitpix == 16 and bs == 1.0 and bz == 32768.0:
                        data = data.astype(np.uint16)
                    else:
                        data = data.astype(np.float32)
                    # the scaling is applied: the cards must not travel with the physical values (a header
                    # copied onto another product would scale it a second time when read back)
                    hdr.remove("BSCALE")
                    hdr.remove("BZERO")
                else:
                    data = data.astype(data.dtype.newbyteorder("="))
        off += ((nbytes + BLOCK - 1) // BLOCK) * BLOCK
        out.append((hdr, data))
    return out


def read_image(path, hdu=0):
    hdr, data = read_hdus(path)[hdu]
    return data, hdr


def read_header(path):
    """The primary header alone (the pixels are not read)."""
    size = 16 * BLOCK
    while True:
        buf = np.fromfile(path, dtype=np.uint8, count=size)
        try:
            hdr = _read_header(buf, 0)[0]
            # as read_hdus: scaling cards other than the neutral ones belong to the stored integers, not to the pixels
            if float(hdr.get("BSCALE", 1.0)) != 1.0 or float(hdr.get("BZERO", 0.0)) != 0.0:
                hdr.remove("BSCALE")
                hdr.remove("BZERO")
            return hdr
        except IOError:
            if len(buf) < size:
                raise
            size *= 4


def _header_bytes(cards80):
    text = "".join(cards80) + "END".ljust(80)
    pad = (-len(text)) % BLOCK
    return (text + " " * pad).encode("ascii")


def write_image(path, data, header=None, extra_hdus=()):
    """Write ``data`` (2-D) as the primary HDU; ``extra_hdus`` are (kind, header, array)."""
    data = np.ascontiguousarray(data)
    bitpix = _BITPIX[data.dtype.name]
    cards = []
    for k, v, c in (("SIMPLE", True, "file does conform to FITS standard"),
                    ("BITPIX", bitpix, "number of bits per data pixel"),
                    ("NAXIS", 2, "number of data axes"),
                    ("NAXIS1", data.shape[1], "length of data axis 1"),
                    ("NAXIS2", data.shape[0], "length of data axis 2")):
        cards += _cards_for(k, v, c)
    skip = {"SIMPLE", "BITPIX", "NAXIS", "NAXIS1", "NAXIS2", "END"}
    unsigned16 = data.dtype == np.uint16          # FITS has no unsigned 16-bit type: int16 + BZERO 32768
    if header is not None:
        for k, v, c in header.cards:
            if k in skip:
                continue
            if k in ("BSCALE", "BZERO"):
                # pixels are written as they are: only the neutral values may stay in the header (hotpants writes
                # BZERO 0 / BSCALE 1 on its difference image, main.c:268-276)
                try:
                    neutral = float(str(v)) == (1.0 if k == "BSCALE" else 0.0)
                except ValueError:
                    neutral = False
                if unsigned16 or not neutral:
                    continue
            cards += _cards_for(k, v, c)
    if unsigned16:
        cards += _cards_for("BSCALE", 1, "") + _cards_for("BZERO", 32768, "offset data range to that of unsigned short")
        data = (data.astype(np.int32) - 32768).astype(np.int16)
    with open(path, "wb") as f:
        f.write(_header_bytes(cards))
        raw = np.ascontiguousarray(data.astype(_DT[bitpix], copy=False))      # big-endian copy, written from its own buffer
        f.write(raw.reshape(-1).view(np.uint8))
        f.write(b"\0" * ((-raw.nbytes) % BLOCK))
        for kind, hdr, arr in extra_hdus:
            assert kind == "bintable"
            arr = np.ascontiguousarray(arr, dtype=np.float64)
            nrows, nf = arr.shape
            c2 = []
            for k, v, c in (("XTENSION", "BINTABLE", "binary table extension"), ("BITPIX", 8, ""),
                            ("NAXIS", 2, ""), ("NAXIS1", 8 * nf, ""), ("NAXIS2", nrows, ""),
                            ("PCOUNT", 0, ""), ("GCOUNT", 1, ""), ("TFIELDS", nf, "")):
                c2 += _cards_for(k, v, c)
            for k, v, c in hdr.cards:
                c2 += _cards_for(k, v, c)
            f.write(_header_bytes(c2))
            raw = arr.astype(">f8").tobytes()
            f.write(raw)
            f.write(b"\0" * ((-len(raw)) % BLOCK))
