"""
TEST INFRASTRUCTURE ONLY -- the CPU oracle for the h5parm -> a-term cube hot path.

Nothing in the product package (``ska_sdp_screen_fitting_b200``) imports, links or
executes anything under ``oracle/``.  Only ``tests/``, ``__graft_entry__.smoke()`` and
the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` use it, and there
only as the checker, never as the thing measured as "ours" or shipped.

Two layers:

* ``oracle.refshim``  -- runs the reference's OWN Python modules, unmodified, from
  ``/root/reference/src`` under a compatibility shim (stub astropy / lsmtool / shapely /
  tables, a duck-typed in-memory H5parm).  Exists only in the build container (the
  reference tree is not present on the GPU box).  Used to pin ``oracle.port`` and to
  generate the golden fixtures in ``tests/golden`` (``oracle/make_golden.py``).
* ``oracle.port``     -- a self-contained numpy/scipy restatement of the same
  algorithms, each function citing the reference file:line it follows.  This is what
  travels to the GPU box.

Parity pinning: the reference ships no golden vectors for this path (its two golden
FITS cubes are missing from the mount and were never compared by a test), so
``oracle.port`` is pinned against outputs of the reference itself run here under
``oracle.refshim`` (fixtures + generating script committed).  Third-party arithmetic
that the reference delegates to libraries absent from this image (wcslib, GEOS,
Pillow 9.0.1, scipy 1.7.3 ``pinv`` cutoff semantics) is restated and is, per
SURVEY.md section 8(c), "parity unpinned" at the bit level; see DESIGN.md.
"""
