"""Minimal single-HDU FITS image I/O (numpy only).

The reference writes / reads its PSF file through astropy.io.fits or pyfits (getpsf.py:400-427,
rdpsf.py:36-45); neither is installed on the B200 image, and the PSF file is a plain primary HDU with
a dozen scalar cards and one 2-D float array -- which is all this module handles (FITS standard 4.0:
80-character cards in 2880-byte blocks, big-endian data, BITPIX -64/-32/8/16/32/64)."""
import numpy as np

BLOCK = 2880
_BITPIX = {np.dtype(">f8"): -64, np.dtype(">f4"): -32, np.dtype("u1"): 8, np.dtype(">i2"): 16, np.dtype(">i4"): 32, np.dtype(">i8"): 64}
_DTYPE = {v: k for k, v in _BITPIX.items()}


def _fmt_value(v):
    if isinstance(v, (bool, np.bool_)):
        return "%20s" % ("T" if v else "F")
    if isinstance(v, (int, np.integer)):
        return "%20d" % int(v)
    if isinstance(v, (float, np.floating)):
        v = float(v)
        if v != v or v in (float("inf"), float("-inf")):
            return "'%-8s'" % repr(v)  # FITS has no NaN/inf literal; stored as a string card
        s = repr(v).upper()
        if "." not in s and "E" not in s:
            s += "."
        return "%20s" % s
    s = str(v).replace("'", "''")
    return "'%-8s'" % s


def _card(key, value=None, comment=None):
    if value is None:
        return ("%-8s%s" % (key, " " * 72))[:80]
    body = "%-8s= %s" % (key[:8].upper(), _fmt_value(value))
    if comment:
        body += " / " + str(comment)
    return ("%-80s" % body)[:80]


def write_image(path, data, cards=()):
    """``cards``: iterable of (KEY, value) or (KEY, value, comment), written after the mandatory keywords."""
    a = np.asarray(data)
    if a.dtype.kind == "f":
        be = a.astype(">f8" if a.dtype.itemsize == 8 else ">f4")
    elif a.dtype == np.uint8:
        be = a
    else:
        be = a.astype(">i%d" % max(a.dtype.itemsize, 2))
    head = [_card("SIMPLE", True, "conforms to FITS standard"), _card("BITPIX", _BITPIX[be.dtype], "array data type"),
            _card("NAXIS", a.ndim, "number of array dimensions")]
    for i, n in enumerate(reversed(a.shape)):
        head.append(_card("NAXIS%d" % (i + 1), int(n)))
    for c in cards:
        head.append(_card(*c))
    head.append("%-80s" % "END")
    hbytes = "".join(head).encode("ascii")
    hbytes += b" " * (-len(hbytes) % BLOCK)
    dbytes = be.tobytes()
    dbytes += b"\0" * (-len(dbytes) % BLOCK)
    with open(path, "wb") as fh:
        fh.write(hbytes)
        fh.write(dbytes)


def _parse_value(text):
    t = text.strip()
    if t.startswith("'"):
        end = 1
        while True:  # closing quote that is not a doubled quote
            end = t.index("'", end)
            if t[end:end + 2] == "''":
                end += 2
                continue
            break
        s = t[1:end].replace("''", "'").rstrip()
        if s in ("nan", "inf", "-inf"):
            return float(s)
        return s
    t = t.split("/")[0].strip()
    if t == "T":
        return True
    if t == "F":
        return False
    try:
        return int(t)
    except ValueError:
        return float(t.replace("D", "E"))


def read_image(path):
    """Returns ``(data, header)``; ``header`` is an ordered dict KEY -> value of the primary HDU."""
    with open(path, "rb") as fh:
        raw = fh.read()
    header, pos, done = {}, 0, False
    while not done:
        block = raw[pos:pos + BLOCK].decode("ascii")
        if len(block) < BLOCK:
            raise ValueError("%s: truncated FITS header" % path)
        pos += BLOCK
        for i in range(0, BLOCK, 80):
            card = block[i:i + 80]
            key = card[:8].strip()
            if key == "END":
                done = True
                break
            if card[8:10] == "= ":
                header[key] = _parse_value(card[10:])
    naxis = int(header.get("NAXIS", 0))
    if naxis == 0:
        return None, header
    shape = tuple(int(header["NAXIS%d" % (i + 1)]) for i in reversed(range(naxis)))
    dt = _DTYPE[int(header["BITPIX"])]
    n = int(np.prod(shape))
    data = np.frombuffer(raw, dtype=dt, count=n, offset=pos).reshape(shape)
    scale, zero = header.get("BSCALE", 1), header.get("BZERO", 0)
    out = data.astype(dt.newbyteorder("="))
    if scale != 1 or zero != 0:
        out = out * scale + zero
    return out, header
