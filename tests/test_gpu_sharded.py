"""GPU parity test of config 5's N>1 path: brute-force top-2 over a row-sharded database with the exchange fused into the
search kernel (orbb200_match_sharded_top2_device: peer stores through CUDA-IPC mapped memory + release/acquire flags).
One process per rank; ranks use distinct GPUs when the box has them, else they share cuda:0 (the peers' memory is then mapped
through the same IPC path, and the GPU time-slices the two contexts).  The merged result must equal the flat search of the
CPU oracle on every rank, for several steps in a row (double-buffered gather slots, epoch flags)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = sys.argv[1]
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import orbb200, ref_cpu, synth
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
ngpu = torch.cuda.device_count()
dev = torch.device("cuda", rank % ngpu)
torch.cuda.set_device(dev)
dist.init_process_group("gloo", rank=rank, world_size=world)
nq, ndb = int(sys.argv[2]), int(sys.argv[3])
sm = orbb200.ShardedMatcher(rank, world, nq, dev.index)
ok = True
for step in range(4):
    q, db, _ = synth.descriptor_db(nq, ndb, seed=100 + step)
    db[ndb - 3] = db[17]          # duplicate rows on different shards: the lower global index must win
    q[5] = db[17]
    lo, hi = orbb200.shard_range(ndb, world, rank)
    q_d = torch.from_numpy(q).to(dev)
    db_d = torch.from_numpy(np.ascontiguousarray(db[lo:hi])).to(dev)
    out_d = torch.zeros((nq, 4), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    sm.match_device(q_d.data_ptr(), nq, db_d.data_ptr(), hi - lo, lo, out_d.data_ptr())
    sm.status()
    r = out_d.cpu().numpy()
    flat = ref_cpu.bf_top2(q, db)
    ok = ok and all(np.array_equal(r[:, i], flat[f]) for i, f in enumerate(("best_dist", "best_idx", "second_dist", "second_idx")))
    ok = ok and r[5, 1] == 17
res = torch.tensor([1 if ok else 0])
dist.all_reduce(res, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
print("SHARDED_OK" if int(res) == 1 else "SHARDED_MISMATCH rank %d" % rank)
'''


@pytest.mark.parametrize("world,nq,ndb", [(2, 2048, 40000), (3, 700, 9001)])
def test_sharded_top2_fused_exchange_equals_flat(tmp_path, world, nq, ndb):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = 29700 + os.getpid() % 200 + world
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT, str(nq), str(ndb)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for k in procs:
                k.kill()
            raise
        outs.append(o)
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert all("SHARDED_OK" in o for o in outs), "\n".join(outs)
