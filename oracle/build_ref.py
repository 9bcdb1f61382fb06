"""
TEST INFRASTRUCTURE (oracle): make the reference's own modules available where
``/root/reference`` is not mounted (the GPU box).

The reference is pure Python: the hot path lives in six files plus three small helpers.  This
recipe copies them, unmodified, from ``/root/reference/src`` into ``oracle/_ref/`` -- a
directory that is git-ignored (reference sources never enter the history) but travels to the
GPU box with the tree, exactly like the built ``libssf_b200.so``.  ``oracle.refshim`` imports
the reference from ``/root/reference/src`` when it is mounted and from ``oracle/_ref`` otherwise,
so ``bench.py --impl reference`` and the ``cpu_baseline`` leg time the REFERENCE (``kind:
"reference"``) on the box's host cores; the loop-faithful port (``oracle/port/refloops.py``) is
only the fallback when neither exists.

``utils/h5parm.py`` (pytables object model) is not copied: pytables is not installed anywhere
here and the shim replaces that module (``refshim/fake_h5parm.py``).

    python -m oracle.build_ref          # called by __graft_entry__.build()
"""
import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCE = "/root/reference/src/ska_sdp_screen_fitting"
DEST_ROOT = os.path.join(HERE, "_ref")
DEST = os.path.join(DEST_ROOT, "ska_sdp_screen_fitting")
FILES = ["stationscreen.py", "kl_screen.py", "voronoi_screen.py", "screen.py", "make_aterm_images.py",
         "_version.py", "utils/processing_utils.py", "utils/multiprocmanager.py", "utils/_logging.py",
         "utils/deprecated_alias.py"]


def build():
    """Copies the files if the reference is mounted.  Returns the import root holding the
    reference package (``oracle/_ref``), or None when there is neither a mount nor a copy."""
    if not os.path.isdir(SOURCE):
        return DEST_ROOT if os.path.isfile(os.path.join(DEST, "stationscreen.py")) else None
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SOURCE, rel), os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(src, "rb") as fh:
            data = fh.read()
        manifest[rel] = hashlib.sha256(data).hexdigest()
        if not os.path.exists(dst) or open(dst, "rb").read() != data:
            shutil.copyfile(src, dst)
            os.chmod(dst, 0o644)
    with open(os.path.join(DEST_ROOT, "MANIFEST.json"), "w", encoding="utf8") as fh:
        json.dump({"source": SOURCE, "sha256": manifest}, fh, indent=1)
    return DEST_ROOT


if __name__ == "__main__":
    print(build())
