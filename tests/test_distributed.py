"""world_size-2 ``gloo`` test of the multi-rank path on CPU: time slabs shard across ranks with
no data-path collective; every rank writes its own byte range of the shared FITS cubes.  The
result must be byte-identical to a single-rank run."""
import os
import socket
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, emu_library, golden_dir, out_dir):
    port = str(_free_port())
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "dist_worker.py"), str(r), str(world), port,
                               emu_library, golden_dir, str(out_dir)],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    for p in procs:
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0, out[-3000:]


def test_two_ranks_write_the_same_cubes_as_one(emu_library, golden_dir, tmp_path):
    one, two = tmp_path / "one", tmp_path / "two"
    one.mkdir()
    two.mkdir()
    _run(1, emu_library, golden_dir, one)
    _run(2, emu_library, golden_dir, two)
    names = sorted(os.listdir(one))
    assert names == sorted(os.listdir(two))
    assert {"vor_0.fits", "vor_1.fits", "vor.txt", "kl_0.fits", "kl_1.fits", "kl.txt"} <= set(names)
    for name in names:
        with open(one / name, "rb") as fa, open(two / name, "rb") as fb:
            a, b = fa.read(), fb.read()
        if name.endswith(".txt"):
            assert a.replace(str(one).encode(), b"") == b.replace(str(two).encode(), b"")
        else:
            assert a == b, name


def test_split_range_covers_everything():
    from ska_sdp_screen_fitting_b200.screen import split_range

    for n in (0, 1, 5, 17):
        for parts in (1, 2, 3, 8):
            got = [split_range(10, 10 + n, parts, i) for i in range(parts)]
            assert got[0][0] == 10 and got[-1][1] == 10 + n
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
            assert max(hi - lo for lo, hi in got) - min(hi - lo for lo, hi in got) <= 1
