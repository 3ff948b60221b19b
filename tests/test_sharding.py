"""Multi-GPU sharding: images are independent, so N GPUs = N shards and no collective.
Host logic only: gjd_partition through the C ABI, and a world_size-2 `gloo` run in which every rank
computes the same partition, decodes its own shard (on the CPU via tests/hostsim, the product's
host+device functions stepped serially) and the gather is pointer bookkeeping checked on rank 0."""
import hashlib
import os
import subprocess
import sys

import numpy as np

from conftest import ROOT, golden_files


def test_partition_properties():
    import gjd
    rng = np.random.default_rng(2024)
    sizes = rng.choice([256, 384, 512, 640, 768, 1024, 1536, 2048], 500)
    w = (sizes.astype(np.uint64) ** 2)
    w[::37] = 0                                    # rejected headers take no part
    L = gjd.lib()
    for parts in (1, 2, 4, 8):
        part_of = np.full(len(w), -7, np.int32)
        load = np.zeros(parts, np.uint64)
        assert L.gjd_partition(w.ctypes.data, len(w), parts, part_of.ctypes.data, load.ctypes.data) == 0
        assert np.all(part_of[w == 0] == -1)
        assert np.all((part_of[w > 0] >= 0) & (part_of[w > 0] < parts))
        for p in range(parts):
            assert int(load[p]) == int(w[part_of == p].sum())
        assert int(load.sum()) == int(w.sum())
        # greedy LPT: the spread between parts never exceeds the heaviest item
        assert int(load.max() - load.min()) <= int(w.max())
        again = np.empty_like(part_of)
        L.gjd_partition(w.ctypes.data, len(w), parts, again.ctypes.data, None)
        assert np.array_equal(part_of, again)
    assert L.gjd_partition(w.ctypes.data, len(w), 0, part_of.ctypes.data, None) == gjd.E_ARG
    assert L.gjd_partition(None, 0, 3, None, None) == 0


def test_partition_weights_from_headers(golden):
    import gjd
    names = sorted(golden)
    files = [golden[n] for n in names] + [b"\xff\xd8\xff\xd9", b""]
    part_of, load, weight = gjd.partition(files, 2)
    assert weight[names.index("images/8_401x363")] == 408 * 368
    assert weight[names.index("synth/s10_1x1_q90_o0")] == 64
    assert weight[-1] == 0 and weight[-2] == 0 and part_of[-1] == -1 and part_of[-2] == -1
    assert int(load.sum()) == int(weight.sum())



def test_partition_costs_add_the_byte_term(golden):
    """gjd_partition_costs = padded pixels + byte_weight * file length, 0 for rejected headers; on a set whose
    bytes per pixel vary (two qualities of one size) the byte term is what balances the scan bytes."""
    import gjd
    names = sorted(golden)
    files = [golden[n] for n in names] + [b"\xff\xd8\xff\xd9", b""]
    _, _, weight = gjd.partition(files, 2)
    part_of, load, cost = gjd.partition(files, 2, gjd.PARTITION_BYTE_WEIGHT)
    lens = np.array([len(f) for f in files], np.uint64)
    expect = np.where(weight > 0, weight + np.uint64(gjd.PARTITION_BYTE_WEIGHT) * lens, 0).astype(np.uint64)
    assert np.array_equal(cost, expect)
    assert cost[-1] == 0 and cost[-2] == 0 and part_of[-1] == -1
    assert int(load.sum()) == int(cost.sum())
    L = gjd.lib()
    assert L.gjd_partition_costs(None, 1, 5, None) == gjd.E_ARG and L.gjd_partition_costs(None, 0, 5, None) == 0
    # light and heavy files of one size, the heavy ones first in every tie: pixels alone send them all one way
    small, big = golden["synth/s10_1x1_q90_o0"], golden["synth/s10_1x1_q90_o0"] + bytes(4096)
    mixed = [big, small] * 8
    by_px, _, _ = gjd.partition(mixed, 2)
    by_cost, _, _ = gjd.partition(mixed, 2, gjd.PARTITION_BYTE_WEIGHT)
    spread = lambda po: abs(sum(len(f) for f, p in zip(mixed, po) if p == 0) - sum(len(f) for f, p in zip(mixed, po) if p == 1))
    assert spread(by_cost) == 0 and spread(by_px) > 0


_WORKER = r'''
import hashlib, os, sys
sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np
import torch.distributed as dist
import conftest, gjd
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
g = conftest.golden_files()
names = sorted(g)
files = [g[n] for n in names]
mine, load = gjd.shard_indices(files, rank, world, gjd.PARTITION_BYTE_WEIGHT)      # the weights bench.py --config c5 uses
hs = conftest.HostSim()
res = {}
for i in mine:
    rc, coef, rgb, st = hs.decode(files[i])
    assert rc == 0
    res[names[i]] = hashlib.sha256(rgb.tobytes()).hexdigest()
out = [None] * world
dist.all_gather_object(out, (mine, [int(x) for x in load], res))
dist.barrier()
if rank == 0:
    import json
    print("RESULT " + json.dumps(out))
dist.destroy_process_group()
'''


def test_two_rank_gloo_shards_cover_and_match(ref_outputs, hostsim, tmp_path):
    """world_size 2 over gloo: shards are disjoint, cover the set, agree on the loads, and every image a
    rank decoded equals the compiled reference's pixels."""
    import json
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29500 + os.getpid() % 2000), WORLD_SIZE="2",
               PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "gpu-jpeg-decoder_b200"), os.path.join(ROOT, "tools")]))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-2000:]
    line = [l for l in outs[0][0].splitlines() if l.startswith("RESULT ")][0]
    gathered = json.loads(line[7:])
    names = sorted(golden_files())
    a, b = set(gathered[0][0]), set(gathered[1][0])
    assert not (a & b) and (a | b) == set(range(len(names)))
    assert gathered[0][1] == gathered[1][1]
    merged = dict(gathered[0][2], **gathered[1][2])
    checked = 0
    for n in names:
        want = ref_outputs[n + "/rgb"]
        assert merged[n] == hashlib.sha256(np.ascontiguousarray(want, np.uint8).tobytes()).hexdigest(), n
        checked += 1
    assert checked == len(names)
