"""
Patch positions of a makesourcedb-format sky model (what the reference obtains through
``lsmtool.load(f).getPatchPositions()``, voronoi_screen.py:241-247, kl_screen.py:75-81).

Patch definition lines have empty Name and Type fields:  `` , , Patch_0, 8:37:42.95, 65.13.47.49``
with RA as h:m:s and Dec as d.m.s (resources/skymodel.txt:1-13).
"""


def _sexagesimal(fields):
    head = fields[0].strip()
    sign = -1.0 if head.startswith("-") else 1.0
    vals = [abs(float(head))] + [float(f) for f in fields[1:3]]
    while len(vals) < 3:
        vals.append(0.0)
    return sign * (vals[0] + vals[1] / 60.0 + vals[2] / 3600.0)


def ra_to_deg(text):
    text = text.strip()
    if ":" not in text:
        return float(text)
    return _sexagesimal(text.split(":")) * 15.0


def dec_to_deg(text):
    text = text.strip()
    if ":" in text:
        return _sexagesimal(text.split(":"))
    parts = text.split(".")
    if len(parts) < 3:
        return float(text)
    seconds = parts[2] if len(parts) == 3 else parts[2] + "." + "".join(parts[3:])
    return _sexagesimal([parts[0], parts[1], seconds])


def patch_positions(filename):
    """Returns {patch name: (ra_deg, dec_deg)} in file order."""
    out = {}
    columns = None
    with open(filename, encoding="utf8") as handle:
        for line in handle:
            stripped = line.strip()
            if not stripped or stripped.startswith("#"):
                continue
            if stripped.upper().startswith("FORMAT"):
                spec = stripped.split("=", 1)[1]
                columns = [c.split("=")[0].strip().lower() for c in spec.split(",")]
                continue
            cells = [c.strip() for c in line.split(",")]
            cols = columns or ["name", "type", "patch", "ra", "dec"]
            try:
                i_name, i_type, i_patch = cols.index("name"), cols.index("type"), cols.index("patch")
                i_ra, i_dec = cols.index("ra"), cols.index("dec")
            except ValueError as exc:
                raise ValueError(f"{filename}: sky model has no patch columns") from exc
            if len(cells) <= max(i_patch, i_ra, i_dec):
                continue
            if cells[i_name] == "" and cells[i_type] == "" and cells[i_patch]:
                out[cells[i_patch]] = (ra_to_deg(cells[i_ra]), dec_to_deg(cells[i_dec]))
    if not out:
        raise ValueError(f"{filename}: no patch definitions found")
    return out
