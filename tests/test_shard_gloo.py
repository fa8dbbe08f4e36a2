"""CPU suite, part 3: the N > 1 path.  Textures shard over ranks with no data-path collective
(SURVEY.md 8e); torch.distributed is only used the way bench.py uses it (barrier + MAX of the
elapsed time).  world_size 2 over gloo; the per-texture work is done by the CPU oracle here (this is
a test: the oracle plays the device), and the union of the ranks' outputs must equal a 1-rank run."""
import hashlib
import os
import socket

import numpy as np
import pytest

from rosy_b200 import shard


def test_plan_is_a_deterministic_partition():
    costs = [4096 * 4096] * 10 + [2048 * 2048] * 30 + [1024 * 1024] * 30  # config 3's 70 textures
    for world in (1, 2, 4, 8):
        p = shard.plan(costs, world)
        assert sorted(i for s in p for i in s) == list(range(len(costs)))
        assert p == shard.plan(list(costs), world)
        mx, ratio = shard.imbalance(costs, world)
        assert ratio < 1.06, (world, ratio)
        for r in range(world):
            assert shard.my_items(costs, r, world) == p[r]
    assert shard.plan([], 4) == [[], [], [], []]
    assert shard.plan([5], 2) == [[0], []]
    with pytest.raises(ValueError):
        shard.plan([1], 0)
    with pytest.raises(ValueError):
        shard.my_items([1], 2, 2)


def test_dedupe_fixes_the_reference_tuple_compare():
    assert shard.dedupe([3, 1, 3, 2, 1, 3]) == [3, 1, 2]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _texture_set():
    from rosy_b200 import synth

    sizes = [(64, 64), (32, 32), (64, 32), (16, 16), (32, 64), (48, 48), (8, 8)]
    kinds = ["albedo", "normal", "metal_rough"]
    return [(kinds[i % 3], w, h, i) for i, (w, h) in enumerate(sizes)], synth


def _encode(item, synth):
    import oracle as O

    kind, w, h, idx = item
    img = synth.make(kind, w, h, idx)
    dds = O.generate_texture(img, O.KIND_NORMAL if kind == "normal" else O.KIND_SRGB, O.FMT_BC7, O.Q_FAST, threads=1)
    return hashlib.sha256(dds.tobytes()).hexdigest()


def _worker(rank, world, port, q):
    import time

    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    items, synth = _texture_set()
    mine = shard.my_items([w * h for (_, w, h, _) in items], rank, world)
    dist.barrier()
    t0 = time.perf_counter()
    result = {i: _encode(items[i], synth) for i in mine}
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
    dist.barrier()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # bench.py: max over ranks
    gathered = [None] * world
    dist.all_gather_object(gathered, result)  # test-only: collect digests for the comparison
    if rank == 0:
        q.put((gathered, float(t.item())))
    dist.destroy_process_group()


def test_world_size_2_gloo_covers_every_texture_once_with_identical_bytes():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, tmax = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    items, synth = _texture_set()
    seen = {}
    for part in gathered:
        for i, digest in part.items():
            assert i not in seen, f"texture {i} encoded by two ranks"
            seen[i] = digest
    assert sorted(seen) == list(range(len(items)))
    for i, it in enumerate(items):  # the bytes do not depend on which rank encoded the texture
        assert seen[i] == _encode(it, synth)
    assert tmax > 0.0
