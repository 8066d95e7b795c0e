"""The micro workloads of bench.py at reduced sizes: each prints the contract's JSON line, and the parity of the run
itself against the reference's own routines (oracle/_ref/novum_ref, when it travelled to the box) is part of that line."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.exists(os.path.join(REPO, "oracle", "_ref", "novum_ref"))


def _bench(*args):
    out = subprocess.check_output([sys.executable, os.path.join(REPO, "bench.py")] + list(args), timeout=900).decode()
    line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "e2e", "roofline", "config"):
        assert k in line, k
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and 0 < line["roofline"]["frac"]
    return line


def test_trace_batch_line_and_parity(built):
    line = _bench("--workload", "trace_batch", "--rays-side", "512", "--steps", "2", "--warmup", "1", "--detail", "8")
    assert set(line["sets"]) == {"primary", "secondary", "shadow"}
    if HAVE_REF:
        p = line["parity_vs_reference_routines"]
        for name in ("primary", "secondary"):
            assert p[name]["mismatch"] == 0 and p[name]["exact"] + p[name]["tie"] + p[name]["superset"] == p[name]["n"]
            assert p[name]["tie"] + p[name]["superset"] <= 0.01 * p[name]["n"]
        assert p["shadow"]["bit_equal"] == p["shadow"]["n"]
        assert line["cpu_baseline"]["kind"] == "reference"


def test_hashgrid_line_and_parity(built):
    line = _bench("--workload", "hashgrid", "--grid-small", "--steps", "2", "--warmup", "1")
    c = line["cases"]["small"]
    assert c["photons"] == 200_000 and c["build_ms"] > 0
    if HAVE_REF:
        assert all(c["parity"].values()), c["parity"]


def test_frame_finish_line(built):
    line = _bench("--workload", "frame_finish", "--steps", "5", "--warmup", "2")
    assert line["parity_ok"] is True and line["unit"] == "Mpixels/s"
    assert line["roofline"]["kernel"] == "frameFinishKernel"
