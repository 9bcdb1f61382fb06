"""
TEST INFRASTRUCTURE (oracle): stand-ins for the parts of ``astropy`` the reference uses.

astropy is not installed in this image.  The reference calls it at
  * stationscreen.py:275-300, 226-229, 267-270   (2-axis TAN WCS, crpix 1000, cdelt 0.0005)
  * kl_screen.py:240-254                         (2-axis TAN WCS of the image grid)
  * voronoi_screen.py:234-282                    (6-axis WCS built from the cube header, SIN)
  * processing_utils.py:202-291, screen.py:194,381-382 (astropy.io.fits PrimaryHDU/HDUList)
  * screen.py:12,71-74                           (astropy.coordinates.Angle for str RA/Dec)

The projections are the closed-form gnomonic (TAN) / orthographic (SIN) zenithal
formulas of the FITS WCS paper II (what wcslib evaluates); see SURVEY.md appendix A.
Bit-level equality with wcslib is not claimed ("parity unpinned" for this dependency).
"""
import builtins
import math
import os

import numpy as np

_D2R = math.pi / 180.0
_R2D = 180.0 / math.pi


# ----------------------------------------------------------------------------- wcs
class _WcsParams:
    def __init__(self, naxis):
        self.naxis = naxis
        self.crpix = [0.0] * naxis
        self.cdelt = [1.0] * naxis
        self.crval = [0.0] * naxis
        self.ctype = [""] * naxis
        self.pv = []

    def set_pv(self, pv):
        # PV2_1 is inert for TAN / SIN without obliqueness terms
        self.pv = list(pv)


class WCS:
    """Minimal stand-in for ``astropy.wcs.WCS`` (celestial pair TAN|SIN + linear axes)."""

    def __init__(self, header=None, naxis=None):
        if header is not None:
            naxis = int(header["NAXIS"])
            self.wcs = _WcsParams(naxis)
            for i in range(naxis):
                k = i + 1
                self.wcs.crpix[i] = float(header.get(f"CRPIX{k}", 0.0))
                self.wcs.cdelt[i] = float(header.get(f"CDELT{k}", 1.0))
                self.wcs.crval[i] = float(header.get(f"CRVAL{k}", 0.0))
                self.wcs.ctype[i] = str(header.get(f"CTYPE{k}", ""))
        else:
            self.wcs = _WcsParams(int(naxis))

    @property
    def axis_type_names(self):
        return [c.split("-")[0] for c in self.wcs.ctype]

    def _celestial(self):
        names = self.axis_type_names
        lon = names.index("RA")
        lat = names.index("DEC")
        proj = self.wcs.ctype[lon][-3:]
        return lon, lat, proj

    def wcs_world2pix(self, world, origin):
        world = np.asarray(world, dtype=np.float64)
        out = np.empty_like(world)
        lon, lat, proj = self._celestial()
        w = self.wcs
        for i in range(w.naxis):
            if i in (lon, lat):
                continue
            out[:, i] = (world[:, i] - w.crval[i]) / w.cdelt[i] + w.crpix[i] - (1 - origin)
        # wcslib stores CRVAL/CDELT/CRPIX as C doubles whatever the caller assigned
        a0 = float(w.crval[lon]) * _D2R
        d0 = float(w.crval[lat]) * _D2R
        for r in range(world.shape[0]):
            a = world[r, lon] * _D2R
            d = world[r, lat] * _D2R
            da = a - a0
            xi = math.cos(d) * math.sin(da)
            eta = math.sin(d) * math.cos(d0) - math.cos(d) * math.sin(d0) * math.cos(da)
            if proj == "TAN":
                den = math.sin(d) * math.sin(d0) + math.cos(d) * math.cos(d0) * math.cos(da)
                xi /= den
                eta /= den
            elif proj != "SIN":
                raise NotImplementedError(proj)
            out[r, lon] = _R2D * xi / float(w.cdelt[lon]) + float(w.crpix[lon]) - (1 - origin)
            out[r, lat] = _R2D * eta / float(w.cdelt[lat]) + float(w.crpix[lat]) - (1 - origin)
        return out

    def wcs_pix2world(self, pix, origin):
        pix = np.asarray(pix, dtype=np.float64)
        out = np.empty_like(pix)
        lon, lat, proj = self._celestial()
        w = self.wcs
        for i in range(w.naxis):
            if i in (lon, lat):
                continue
            out[:, i] = (pix[:, i] + (1 - origin) - w.crpix[i]) * w.cdelt[i] + w.crval[i]
        # wcslib stores CRVAL/CDELT/CRPIX as C doubles whatever the caller assigned
        a0 = float(w.crval[lon]) * _D2R
        d0 = float(w.crval[lat]) * _D2R
        for r in range(pix.shape[0]):
            xi = float(w.cdelt[lon]) * (pix[r, lon] + (1 - origin) - float(w.crpix[lon])) * _D2R
            eta = float(w.cdelt[lat]) * (pix[r, lat] + (1 - origin) - float(w.crpix[lat])) * _D2R
            if proj == "TAN":
                den = math.cos(d0) - eta * math.sin(d0)
                a = a0 + math.atan2(xi, den)
                d = math.atan2(math.sin(d0) + eta * math.cos(d0), math.hypot(xi, den))
            elif proj == "SIN":
                zeta = math.sqrt(max(0.0, 1.0 - xi * xi - eta * eta))
                d = math.asin(eta * math.cos(d0) + zeta * math.sin(d0))
                a = a0 + math.atan2(xi, zeta * math.cos(d0) - eta * math.sin(d0))
            else:
                raise NotImplementedError(proj)
            ra = a * _R2D
            # wcslib normalises celestial longitude to [0, 360)
            ra = ra % 360.0
            out[r, lon] = ra
            out[r, lat] = d * _R2D
        return out


# ----------------------------------------------------------------------- coordinates
class _Quantity:
    def __init__(self, value):
        self.value = value


class Angle:
    """Only what screen.py:71-74 needs: ``Angle("12.5d").to("deg").value``."""

    def __init__(self, text):
        t = str(text).strip().lower()
        if t.endswith("deg"):
            self._deg = float(t[:-3])
        elif t.endswith("d"):
            self._deg = float(t[:-1])
        elif t.endswith("rad"):
            self._deg = float(t[:-3]) * _R2D
        else:
            self._deg = float(t)

    def to(self, unit):
        assert unit == "deg"
        return _Quantity(self._deg)


# ------------------------------------------------------------------------------ fits
class Header:
    """Ordered card list with dict-style access."""

    def __init__(self):
        self._keys = []
        self._vals = {}

    def __setitem__(self, key, val):
        if key not in self._vals:
            self._keys.append(key)
        self._vals[key] = val

    def __getitem__(self, key):
        return self._vals[key]

    def __contains__(self, key):
        return key in self._vals

    def get(self, key, default=None):
        return self._vals.get(key, default)

    def keys(self):
        return list(self._keys)

    def items(self):
        return [(k, self._vals[k]) for k in self._keys]


def _card(key, val):
    if isinstance(val, bool):
        v = f"{'T' if val else 'F':>20}"
    elif isinstance(val, (int, np.integer)):
        v = f"{int(val):>20d}"
    elif isinstance(val, (float, np.floating)):
        s = repr(float(val)).upper()
        if "E" not in s and "." not in s and "N" not in s:
            s += ".0"
        v = f"{s:>20}"
    else:
        s = str(val).replace("'", "''")
        v = f"'{s:<8}'"
        v = f"{v:<20}"
    return f"{key:<8}= {v}".ljust(80)[:80]


def _parse_card(card):
    key = card[:8].strip()
    if card[8:10] != "= ":
        return key, None
    body = card[10:]
    if body.lstrip().startswith("'"):
        s = body.lstrip()[1:]
        end = s.find("'")
        while end != -1 and s[end:end + 2] == "''":
            end = s.find("'", end + 2)
        return key, s[:end].replace("''", "'").rstrip()
    tok = body.split("/")[0].strip()
    if tok in ("T", "F"):
        return key, tok == "T"
    try:
        return key, int(tok)
    except ValueError:
        return key, float(tok.replace("D", "E"))


class PrimaryHDU:
    def __init__(self, data=None, header=None):
        self.data = data
        self.header = header if header is not None else Header()


class HDUList(list):
    def __init__(self, hdus=()):
        super().__init__(hdus)

    def writeto(self, name, overwrite=False):
        if os.path.exists(name) and not overwrite:
            raise OSError(f"{name} exists")
        hdu = self[0]
        data = np.asarray(hdu.data)
        bitpix = {np.dtype("float32"): -32, np.dtype("float64"): -64}[data.dtype]
        cards = [_card("SIMPLE", True), _card("BITPIX", bitpix), _card("NAXIS", data.ndim)]
        for i, n in enumerate(data.shape[::-1]):
            cards.append(_card(f"NAXIS{i + 1}", n))
        cards.append(_card("EXTEND", True))
        skip = {"SIMPLE", "BITPIX", "NAXIS", "EXTEND"} | {
            f"NAXIS{i + 1}" for i in range(data.ndim)
        }
        for k, v in hdu.header.items():
            if k not in skip:
                cards.append(_card(k, v))
        cards.append("END".ljust(80))
        text = "".join(cards)
        text += " " * ((-len(text)) % 2880)
        raw = data.astype(data.dtype.newbyteorder(">")).tobytes()
        with builtins.open(name, "wb") as f:
            f.write(text.encode("ascii"))
            f.write(raw)
            f.write(b"\0" * ((-len(raw)) % 2880))

    def close(self):
        pass


def open(name, memmap=False):  # noqa: A001 - mirrors astropy.io.fits.open
    del memmap
    with builtins.open(name, "rb") as f:
        blob = f.read()
    header = Header()
    pos = 0
    done = False
    while not done:
        block = blob[pos:pos + 2880].decode("ascii")
        pos += 2880
        for c in range(36):
            card = block[c * 80:(c + 1) * 80]
            if card.startswith("END"):
                done = True
                break
            key, val = _parse_card(card)
            if key and val is not None:
                header[key] = val
    naxis = header["NAXIS"]
    shape = tuple(header[f"NAXIS{i}"] for i in range(naxis, 0, -1))
    dt = {-32: ">f4", -64: ">f8"}[header["BITPIX"]]
    data = np.frombuffer(blob, dtype=dt, count=int(np.prod(shape)), offset=pos)
    data = data.reshape(shape).astype(dt[1:])  # native, writable copy
    return HDUList([PrimaryHDU(data, header)])
