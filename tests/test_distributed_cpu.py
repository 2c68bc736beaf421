"""CPU, world_size 2, gloo: the host-side logic of the N>1 paths -- frame / database-row sharding and the
all-gather layout that feeds the top-2 merge.  The per-shard searches are done by the CPU oracle here (checker only);
on GPUs the same gather feeds k_merge_top2 (covered by tests/test_gpu_match.py::test_sharded_merge_equals_flat)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lex_merge(parts):
    """numpy statement of the merge rule: the two lexicographically smallest (dist, global index) per query"""
    world, nq, _ = parts.shape
    out = np.zeros((nq, 4), np.int32)
    for i in range(nq):
        c = []
        for s in range(world):
            d1, i1, d2, i2 = parts[s, i]
            c.append((d1, i1 if i1 >= 0 else 1 << 30))
            c.append((d2, i2 if i2 >= 0 else 1 << 30))
        c.sort()
        (a, ai), (b, bi) = c[0], c[1]
        out[i] = (a, ai if ai < (1 << 30) else -1, b, bi if bi < (1 << 30) else -1)
    return out


def _worker(rank, world, port, ret):
    for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
        sys.path.insert(0, p)
    import orbb200
    import ref_cpu
    import synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nq, ndb = 64, 4099  # not divisible by the world size: ragged shards
    q, db, _ = synth.descriptor_db(nq, ndb, seed=11)
    db[4000] = db[17]  # duplicate rows on different shards: the lower global index must win
    q[5] = db[17]
    lo, hi = orbb200.shard_range(ndb, world, rank)
    p = ref_cpu.bf_top2(q, db[lo:hi])
    part = np.stack([p["best_dist"], np.where(p["best_idx"] >= 0, p["best_idx"] + lo, -1), p["second_dist"],
                     np.where(p["second_idx"] >= 0, p["second_idx"] + lo, -1)], axis=1).astype(np.int32)
    gathered = orbb200.gather_top2_parts(torch.from_numpy(part)).numpy()
    merged = _lex_merge(gathered)
    flat = ref_cpu.bf_top2(q, db)
    ok = (np.array_equal(merged[:, 0], flat["best_dist"]) and np.array_equal(merged[:, 1], flat["best_idx"]) and
          np.array_equal(merged[:, 2], flat["second_dist"]) and np.array_equal(merged[:, 3], flat["second_idx"]))
    # frame sharding: every frame owned exactly once
    owned = torch.zeros(37, dtype=torch.int32)
    flo, fhi = orbb200.shard_range(37, world, rank)
    owned[flo:fhi] += 1
    dist.all_reduce(owned)
    ok = ok and bool((owned == 1).all()) and merged[5, 1] == 17
    ret[rank] = ok
    dist.destroy_process_group()


def test_sharded_top2_and_frame_sharding_world2():
    world = 2
    port = 29600 + os.getpid() % 300
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_shard_range_partitions():
    sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
    import orbb200
    for n in (0, 1, 7, 8, 1 << 20):
        for world in (1, 2, 3, 8):
            spans = [orbb200.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
