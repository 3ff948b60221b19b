"""bench.py's host logic (no GPU): workload description shared by both arms, resident chunking, the parity sample,
the oracle digests it checks against, and the --impl reference arm's JSON line; plus (-m gpu) the c5 workload at
reduced count through the very code the driver runs."""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, ROOT)


def test_config_dict_is_shared_by_both_arms():
    import bench
    for cfg, n in (("c3", 3000), ("c5", 24000)):
        d = bench.config_dict(cfg, n, 1)
        assert d["config"] == cfg and str(n) in d["workload"]
        assert d["scaling"] == ("weak" if cfg == "c3" else "strong")
        assert "step" not in d          # the bounded-sample text lives in cpu_baseline.sample
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": cfg') == 2     # the gjd arm and the reference arm print the same dict


def test_chunks_and_samples():
    import bench
    ids = [i % 7 for i in range(100)]
    weight = {i: 1_000_000_000 + i for i in range(7)}
    chunks = bench.make_chunks(ids, weight)
    assert [p for c in chunks for p in c] == list(range(100))
    assert all(sum(weight[ids[p]] for p in c) <= max(bench.CHUNK_PIXELS, max(weight.values())) for c in chunks)
    s = bench.pick_samples(ids)
    assert sorted(ids[p] for p in s) == list(range(7)) and len(set(s)) == 7
    assert bench.pick_samples([]) == []


def test_oracle_digests_match_golden(golden, ref_outputs):
    """The digests bench.py compares the GPU output with are the unmodified reference's pixels."""
    import bench
    names = sorted(golden)[:6]
    got = bench.oracle_digests([golden[n] for n in names], range(len(names)), 4)
    for i, n in enumerate(names):
        assert got[i] == hashlib.md5(np.ascontiguousarray(ref_outputs[n + "/rgb"]).tobytes()).digest(), n


def test_reference_arm_line():
    import gen_synth
    if not gen_synth.have_cjpeg() and not os.path.isdir(gen_synth.CACHE):
        pytest.skip("no synthetic inputs here")
    env = dict(os.environ, GJD_REF_STEP_SECONDS="0.2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         check=True, stdout=subprocess.PIPE, text=True, env=env).stdout
    r = json.loads(out.strip().splitlines()[-1])
    import bench
    assert r["impl"] == "reference" and r["value"] > 0 and r["e2e"]["value"] == r["value"]
    assert r["config"] == bench.config_dict("c3", 3000, 1)
    assert r["cpu_baseline"]["kind"] in ("reference", "port") and r["cpu_baseline"]["cores"] >= 1


@pytest.mark.gpu
def test_bench_c5_reduced_count():
    """BASELINE config 5 through bench.py itself at 300 images: sharding by gjd_partition, resident chunks,
    all three timed loops parity-checked against the oracle (one image per distinct file)."""
    env = dict(os.environ, GJD_BENCH_CHUNK_GPIX="0.2")          # several resident chunks even at this size
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", "c5", "--images", "300", "--steps", "2",
                          "--warmup", "1", "--e2e-steps", "1", "--no-cpu-baseline"],
                         check=True, stdout=subprocess.PIPE, text=True, env=env).stdout
    r = json.loads(out.strip().splitlines()[-1])
    assert r["config"]["config"] == "c5" and r["scaling"] == "strong"
    assert r["parity_checked"] == 256 and r["parity_detail"]["value"] == 256
    assert r["workload_detail"]["images_decoded_per_step"] == 300 and r["workload_detail"]["resident_chunks_rank0"] > 1
    assert r["value"] > 0 and r["e2e"]["value"] > 0 and r["gpu_launches"] > 0
    assert r["shards"]["imbalance"] == 0.0


@pytest.mark.gpu
def test_bench_c3_reduced_count():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--images", "128", "--steps", "2", "--warmup", "1",
                          "--e2e-steps", "1", "--no-cpu-baseline"], check=True, stdout=subprocess.PIPE, text=True).stdout
    r = json.loads(out.strip().splitlines()[-1])
    assert r["parity_checked"] == 64 and r["config"]["config"] == "c3" and r["roofline"]["frac"] > 0
