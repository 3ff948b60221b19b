"""The oracle (oracle/gjd_oracle.c, a C restatement of the reference cpp-decoder) against every pin
the reference offers: its committed ground-truth arrays, outputs of the unmodified reference compiled
into oracle/_ref (stored in tests/golden/ref_outputs.npz by tools/make_golden.py), and the one
coefficient-level known answer in the reference tree (legacy_versions/cudaU-implementation/gold.txt)."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_files

NAMES = sorted(golden_files())


@pytest.mark.parametrize("name", [n for n in NAMES if n.startswith("images/")])
def test_oracle_matches_reference_ground_truth(oracle, golden, ground_truth, name):
    stem = name.split("/")[1]
    key = stem if stem in ground_truth.files else "regen/" + stem
    rc, _, rgb, _, W, H = oracle.decode(golden[name], want_coef=False)
    assert rc == 0
    assert rgb.shape == ground_truth[key].shape
    assert np.array_equal(rgb, ground_truth[key])


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_compiled_reference(oracle, golden, ref_outputs, name):
    rc, coef, rgb, _, W, H = oracle.decode(golden[name])
    assert rc == 0
    assert np.array_equal(coef, ref_outputs[name + "/coef"]), "quantised coefficients differ"
    assert np.array_equal(rgb, ref_outputs[name + "/rgb"]), "pixels differ"


def test_gold_txt_known_answer(oracle, golden):
    # legacy_versions/cudaU-implementation/gold.txt: first 10 zigzag coefficients of the first Y/Cb/Cr blocks
    gold = [[-268, -68, -8], [-19, 0, 3], [-13, 0, -2], [-7, 0, 1], [0, 0, 2], [-1, 0, 0], [1, 0, 0], [7, 0, 0], [0, 0, 0], [-7, 0, 0]]
    rc, coef, _, _, W, H = oracle.decode(golden["images/1_320x240"], want_rgb=False)
    c = coef.reshape(-1, 3, 64)[0]
    assert [[int(c[0][k]), int(c[1][k]), int(c[2][k])] for k in range(10)] == gold


def test_oracle_array_writer_bytes(oracle, golden, tmp_path):
    sha = json.load(open(os.path.join(GOLDEN, "array_sha256.json")))
    for stem, want in sha.items():
        rc, _, rgb, _, W, H = oracle.decode(golden["images/" + stem], want_coef=False)
        p = str(tmp_path / (stem + ".array"))
        assert oracle.lib.gjdo_write_array(p.encode(), W, H, rgb.ctypes.data) > 0
        assert hashlib.sha256(open(p, "rb").read()).hexdigest() == want


def test_oracle_rejects_truncated(oracle, golden):
    data = golden["images/3_120x120"]
    rc, *_ = oracle.decode(data[: len(data) // 2])
    assert rc != 0
