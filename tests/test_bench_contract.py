"""Host logic of bench.py that needs no GPU: the reference arm's JSON line (bench.py --impl
reference) carries the contract's keys and the SAME `config` as the GPU arm prints for the
workload, and only rank 0 produces it."""
import argparse
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture()
def bench(monkeypatch):
    monkeypatch.syspath_prepend(ROOT)
    monkeypatch.syspath_prepend(os.path.join(ROOT, "tests"))
    mod = importlib.import_module("bench")
    # the timed CPU sample itself is not under test (tens of seconds of the reference's loops)
    monkeypatch.setattr(mod, "cpu_arm", lambda workload, smooth_pix, ncpu: (1.5e6, 2.0, "reference", ncpu, "a bounded sample"))
    return mod


def _args(workload, gpus=1):
    return argparse.Namespace(workload=workload, gpus=gpus, steps=2, warmup=1, slots=0, smooth_pix=-1.0)


@pytest.mark.parametrize("workload, image, directions, columns", [("lofar", "512x512", 45, 5952),
                                                                   ("voronoi1k", "1024x1024", 45, 1488),
                                                                   ("skalow", "2048x2048", 100, 512)])
def test_reference_arm_line(bench, monkeypatch, workload, image, directions, columns):
    monkeypatch.setenv("RANK", "0")
    line = bench.run_reference(_args(workload))
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "impl"):
        assert key in line, key
    assert line["impl"] == "reference" and line["vs_baseline"] is None and line["higher_is_better"] is True
    assert line["metric"] == bench.METRIC and line["unit"] == bench.UNIT
    assert line["value"] == pytest.approx(1.5e6) and line["ms_per_step"] == pytest.approx(2000.0)
    assert line["e2e"] == {"value": line["value"], "unit": bench.UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    base = line["cpu_baseline"]
    assert base["kind"] == "reference" and base["value"] == line["value"] and base["cores"] >= 1 and base["sample"]
    # the GPU arm's config of the same workload: image, directions, columns and slots of ITS step
    cfg = line["config"]
    assert cfg["image"] == image and cfg["directions"] == directions and cfg["columns_per_step"] == columns
    slots = bench.auto_slots(workload)
    assert cfg == bench.workload_config(workload, int(image.split("x")[0]), directions, columns, slots)


def test_reference_arm_runs_on_rank_0_only(bench, monkeypatch):
    monkeypatch.setenv("RANK", "1")
    assert bench.run_reference(_args("lofar", gpus=2)) is None
