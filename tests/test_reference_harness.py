"""The reference's OWN acceptance scripts, unmodified, against gjd-decoder/ (INTEGRATION.md §1).

oracle/Makefile stages testing/compare.py and testing/compare_with_tolerance.py (copied as they are, never into the
repo's history) into oracle/_ref/testing/.  The test lays out a testing/-like directory in tmp — the scripts, images/,
ground_truth/ (the reference's committed .array files, re-emitted from tests/golden/ground_truth.npz and checked against
their sha256) and the output directory the reference's decoders expect to exist — and runs the scripts with
cwd=testing/, exactly as testing/test.sh does, once per image that has a ground truth in the mirror."""
import hashlib
import json
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

STAGED = os.path.join(ROOT, "oracle", "_ref", "testing")


def emit_array(path, rgb):
    """The .array text of JPEGParser::write (cpp-decoder/src/parser.cpp:197-210), written here in Python."""
    _, h, w = rgb.shape
    with open(path, "w") as f:
        f.write("%d %d\n" % (h, w))
        for c in range(3):
            f.write("".join("%d " % v for v in rgb[c].ravel()))
            if c < 2:
                f.write("\n")


def layout(tmp_path, ground_truth):
    sha = json.load(open(os.path.join(GOLDEN, "array_sha256.json")))
    t = tmp_path / "testing"
    (t / "images").mkdir(parents=True)
    (t / "ground_truth").mkdir()
    (t / "gjd_output_arrays").mkdir()          # the reference's decoders do not create theirs either
    for s in ("compare.py", "compare_with_tolerance.py"):
        shutil.copy(os.path.join(STAGED, s), t / s)
    stems = []
    for key in ground_truth.files:
        if "/" in key:          # regen/4_800x600: the mirror has no committed ground truth for it
            continue
        stem = key
        shutil.copy(os.path.join(GOLDEN, "images", stem + ".jpg"), t / "images" / (stem + ".jpg"))
        emit_array(t / "ground_truth" / (stem + ".array"), ground_truth[key])
        assert hashlib.sha256((t / "ground_truth" / (stem + ".array")).read_bytes()).hexdigest() == sha[stem], stem
        stems.append(stem)
    os.symlink(os.path.join(ROOT, "gjd-decoder"), tmp_path / "gjd-decoder")
    return t, stems


def test_ground_truth_arrays_reemitted_exactly(tmp_path, ground_truth):
    """CPU: the ground-truth files the GPU test feeds the reference's scripts are byte-identical to the reference's."""
    if not os.path.exists(os.path.join(STAGED, "compare.py")):
        pytest.skip("oracle/_ref/testing not staged (make -C oracle ref needs /root/reference)")
    t, stems = layout(tmp_path, ground_truth)
    assert len(stems) == 6


@pytest.mark.gpu
@pytest.mark.parametrize("script", ["compare.py", "compare_with_tolerance.py"])
def test_reference_acceptance_scripts(tmp_path, ground_truth, script):
    if not os.path.exists(os.path.join(STAGED, script)):
        pytest.skip("oracle/_ref/testing not staged")
    t, stems = layout(tmp_path, ground_truth)
    env = dict(os.environ)
    try:
        import cv2  # noqa: F401  (compare.py imports it for a preview image only)
    except Exception:
        stub = tmp_path / "stub"
        stub.mkdir()
        (stub / "cv2.py").write_text("def merge(x):\n    return x\n\ndef imwrite(*a, **k):\n    return True\n")
        env["PYTHONPATH"] = str(stub) + os.pathsep + env.get("PYTHONPATH", "")
    for stem in stems:
        r = subprocess.run([sys.executable, script, "../gjd-decoder", "./images/%s.jpg" % stem], cwd=str(t), env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 0, r.stderr
        assert "Congratulations! Output matches ground truth!" in r.stdout, (stem, r.stdout)
        assert "does not match" not in r.stdout
